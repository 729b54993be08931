#!/bin/bash
# Installs the UNMODIFIED reference (/root/reference) into baseline/_ref, offline.
#  * the reference's setup.py lists horovod in setup_requires, which setuptools tries to download; the
#    dist-info stubs in baseline/shim make that requirement resolve locally (PYTHONPATH), nothing is fetched;
#  * /root/reference is read-only and the build writes egg-info into the source tree -> install from a copy;
#  * the reference ships distributed_dot_product/utils/ without __init__.py, so its own find_packages() omits
#    utils/comm.py; the file is copied verbatim next to the installed package.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
rm -rf "$HERE/_ref" /tmp/refsrc
cp -r /root/reference /tmp/refsrc
(cd /tmp/refsrc && PYTHONPATH="$HERE/shim" python -m pip install --no-index --no-build-isolation --no-deps \
    --find-links /opt/wheelhouse --target "$HERE/_ref" /tmp/refsrc)
mkdir -p "$HERE/_ref/distributed_dot_product/utils"
cp /root/reference/distributed_dot_product/utils/comm.py "$HERE/_ref/distributed_dot_product/utils/comm.py"
for f in module.py multiplication/functions.py multiplication/ops.py utils/comm.py; do
  cmp /root/reference/distributed_dot_product/$f "$HERE/_ref/distributed_dot_product/$f"
done
echo "reference installed in $HERE/_ref (byte-identical sources)"
