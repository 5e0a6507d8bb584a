#!/bin/bash
# ./setup.sh [--force]
#
# One-step set-up, the role of the reference's 2-setup-host-and-build-container.sh (which runs install-scripts/setup.sh and
# then build-container.sh, whose last act is `singularity run` as a sanity check: build-container.sh:29-30). There is no
# container, compiler or MPI to install here: the step builds the two native libraries in-tree for sm_100a and prints the
# environment report (utils/doctor.py = the images' %runscript role). Exit code is the report's.
set -e
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
export PYTHONPATH="${HERE}/..:${PYTHONPATH}"
if [ "$1" == "--force" ]; then
    python -m azure_hc_intel_tf_b200.build --force
else
    python -m azure_hc_intel_tf_b200.build
fi
exec python -m azure_hc_intel_tf_b200.utils.doctor
