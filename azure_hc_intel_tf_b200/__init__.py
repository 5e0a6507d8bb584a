"""azure_hc_intel_tf_b200 — a B200-native data-parallel CNN training benchmark framework with the
capabilities of the md-k-sarker/azure-hc-intel-tf harness (Intel-TensorFlow + Horovod + tf_cnn_benchmarks).

Sub-packages: ``cli`` (tf_cnn_benchmarks flag surface), ``hvd`` (Horovod-style API), ``models`` (zoo),
``ops`` (layer ops over the sm_100a kernel library), ``parallel`` (launcher, topology, symmetric memory,
allreduce kernels), ``utils`` (flags, checkpoints, tf variable names, logging, timeline), ``bench``
(OSU-style sweeps and kernel micro-benchmarks).
"""
__version__ = "0.1.0"
