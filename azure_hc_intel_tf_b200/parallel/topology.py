"""Single-box topology discovery: GPUs, NVLink links, peer access, multicast (NVLS) support, NUMA placement.

Role in the reference stack: what `lscpu` (benchmark-scripts/run-tf-sing-ucx-openmpi.sh:37-38), `ibv_devinfo`
(azure-scripts/prep-cluster.sh:23) and the hostfile from `nmap` discovery (azure-scripts/setup-pwdless-ssh.sh:32)
tell the reference's launcher about sockets, fabric health and nodes. On one NVSwitch box the questions are: how
many GPUs, are all NVLink links up, can every GPU reach every peer, which NUMA node feeds each GPU.

    python -m azure_hc_intel_tf_b200.parallel.topology          # prints a report, exit code 1 if the fabric is degraded
"""
from __future__ import annotations

import json
import os
import sys
from dataclasses import asdict, dataclass, field
from typing import Dict, List, Optional


@dataclass
class GpuInfo:
    index: int
    name: str = ""
    pci_bus_id: str = ""
    numa_node: int = -1
    memory_gib: float = 0.0
    nvlink_links_up: int = 0
    nvlink_links_total: int = 0
    multicast_supported: Optional[bool] = None
    peers: List[int] = field(default_factory=list)      # indices this GPU can access directly


@dataclass
class Topology:
    gpus: List[GpuInfo] = field(default_factory=list)
    driver: str = ""
    all_to_all_p2p: bool = False
    nvlink_healthy: bool = False
    notes: List[str] = field(default_factory=list)

    def to_json(self) -> str:
        return json.dumps(asdict(self), indent=1)

    def numa_of(self, gpu: int) -> int:
        return self.gpus[gpu].numa_node if gpu < len(self.gpus) else -1

    def report(self) -> List[str]:
        out = [f"GPUs: {len(self.gpus)}   driver: {self.driver}",
               f"peer access all-to-all: {self.all_to_all_p2p}   NVLink links all up: {self.nvlink_healthy}"]
        for g in self.gpus:
            mc = {True: "yes", False: "no", None: "?"}[g.multicast_supported]
            out.append(f"  gpu{g.index}: {g.name}  {g.memory_gib:.0f} GiB  pci {g.pci_bus_id}  numa {g.numa_node}  "
                       f"nvlink {g.nvlink_links_up}/{g.nvlink_links_total} up  multicast {mc}  peers {g.peers}")
        out += [f"  note: {n}" for n in self.notes]
        return out


def _numa_node(pci_bus_id: str) -> int:
    # NVML reports 00000000:1B:00.0, sysfs uses 0000:1b:00.0
    bid = pci_bus_id.lower()
    if len(bid.split(":")[0]) == 8:
        bid = bid[4:]
    try:
        with open(f"/sys/bus/pci/devices/{bid}/numa_node") as f:
            return int(f.read().strip())
    except Exception:
        return -1


def _multicast_supported(index: int) -> Optional[bool]:
    try:
        from cuda import cuda   # cuda-python
        err, = cuda.cuInit(0)
        err, dev = cuda.cuDeviceGet(index)
        err, val = cuda.cuDeviceGetAttribute(cuda.CUdevice_attribute.CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev)
        return bool(val) if int(err) == 0 else None
    except Exception:
        return None


def discover() -> Topology:
    topo = Topology()
    try:
        import torch
        n = torch.cuda.device_count() if torch.cuda.is_available() else 0
    except Exception:
        torch, n = None, 0
    if n == 0:
        topo.notes.append("no CUDA device visible (CPU plumbing mode)")
        return topo
    nvml = None
    try:
        import pynvml
        pynvml.nvmlInit()
        nvml = pynvml
        topo.driver = pynvml.nvmlSystemGetDriverVersion()
        if isinstance(topo.driver, bytes):
            topo.driver = topo.driver.decode()
    except Exception as e:   # NVML missing: fall back to what torch knows
        topo.notes.append(f"NVML unavailable ({e.__class__.__name__}); link state unknown")
    for i in range(n):
        props = torch.cuda.get_device_properties(i)
        g = GpuInfo(index=i, name=props.name, memory_gib=props.total_memory / 2 ** 30)
        if nvml is not None:
            try:
                h = nvml.nvmlDeviceGetHandleByIndex(i)
                bus = nvml.nvmlDeviceGetPciInfo(h).busId
                g.pci_bus_id = bus.decode() if isinstance(bus, bytes) else bus
                g.numa_node = _numa_node(g.pci_bus_id)
                for link in range(18):
                    try:
                        st = nvml.nvmlDeviceGetNvLinkState(h, link)
                        g.nvlink_links_total += 1
                        g.nvlink_links_up += int(st == 1)
                    except Exception:
                        break
            except Exception as e:
                topo.notes.append(f"gpu{i}: NVML query failed ({e.__class__.__name__})")
        g.multicast_supported = _multicast_supported(i)
        g.peers = [j for j in range(n) if j != i and torch.cuda.can_device_access_peer(i, j)]
        topo.gpus.append(g)
    topo.all_to_all_p2p = all(len(g.peers) == n - 1 for g in topo.gpus)
    topo.nvlink_healthy = all(g.nvlink_links_total == 0 or g.nvlink_links_up == g.nvlink_links_total for g in topo.gpus)
    if n > 1 and not topo.all_to_all_p2p:
        topo.notes.append("some GPU pairs lack peer access: the nvl fabric needs all-to-all P2P; use --fabric=nccl")
    return topo


def main() -> int:
    t = discover()
    print("\n".join(t.report()))
    if "--json" in sys.argv:
        print(t.to_json())
    healthy = (len(t.gpus) <= 1) or (t.all_to_all_p2p and t.nvlink_healthy)
    return 0 if healthy else 1


if __name__ == "__main__":
    sys.exit(main())
