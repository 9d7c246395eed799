#!/bin/sh
# Recipe (TEST/BENCH INFRASTRUCTURE): make the UNMODIFIED reference runnable on the GPU box.
# The reference is ten pure-Python files with no setup.py, so it is not pip-installable; this copies
# them byte for byte from where they lie (/root/reference, present in the build container only)
# into oracle/_ref/, which is git-ignored (never enters history) but NOT gpurun-ignored (it travels
# with the snapshot, like the built libsplitnn.so).  Nothing under splitlearning_b200/ imports it.
# bench.py --impl reference and the cpu_baseline leg run it under torch.distributed.rpc through
# oracle/ref_runner.py; when oracle/_ref is absent they fall back to the torch-CPU port
# (oracle/cpu_baseline.py) and say so ("kind": "port").
set -e
HERE=$(cd "$(dirname "$0")" && pwd)
SRC=${SN_REFERENCE_DIR:-/root/reference}
DST="$HERE/_ref"
if [ ! -f "$SRC/split_nn.py" ]; then
  echo "fetch_ref: $SRC not present; keeping $DST as it is" >&2
  exit 0
fi
rm -rf "$DST"
mkdir -p "$DST/data/mnist_flat"
for f in split_nn.py models.py data_entities.py data_entities_vanilla.py data_entities_vanilla_sisa.py \
         data_entities_sisa_concat.py; do
  cp "$SRC/$f" "$DST/$f"
done
cp "$SRC/data/__init__.py" "$SRC/data/noniid_partition.py" "$SRC/data/relational_table_preprocessor.py" "$DST/data/"
cp "$SRC/data/mnist_flat/__init__.py" "$SRC/data/mnist_flat/mnist_flat_generator.py" "$DST/data/mnist_flat/"
( cd "$SRC" && sha256sum split_nn.py models.py data_entities*.py data/*.py data/mnist_flat/*.py ) > "$DST/SHA256SUMS"
echo "fetch_ref: reference copied to $DST"
