"""Stall inspector of the C++ host runtime (Horovod's stall check; SURVEY.md §5.3 — absent from the reference's
own scripts, provided by the Horovod library it installs)."""
import time

from azure_hc_intel_tf_b200 import _native
from azure_hc_intel_tf_b200.hvd.fusion import FusionPlan


def test_warns_when_a_bucket_stays_in_flight():
    plan = FusionPlan([1000, 1000], [4, 4], [0, 0], [1, 1])
    tr = plan.tracker()
    tr.begin_step()
    assert tr.ready(0) == [] and tr.ready(1) == [0]        # bucket 0 is now "launched" and never marked done
    lib = _native.host()
    st = lib.hvd_stall_start(tr._h, 0.05, 0.1, 0.0)
    time.sleep(0.5)
    warnings = lib.hvd_stall_warnings(st)
    tr.done(0)
    time.sleep(0.2)
    quiet = lib.hvd_stall_warnings(st)
    time.sleep(0.2)
    assert lib.hvd_stall_warnings(st) == quiet               # no new warnings once the bucket completed
    lib.hvd_stall_stop(st)
    assert warnings >= 1
