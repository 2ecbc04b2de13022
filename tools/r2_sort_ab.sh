#!/bin/bash
# A/B of the radix sort at 2^26 (configs[2]): ballot-ranking pass (default) vs the column-counter pass
# (COL_SORT_COLUMN=1), with the result checked against torch.sort(stable=True).   bash tools/r2_sort_ab.sh TAG
TAG=${1:-r2x}
O=gpurun_out; mkdir -p $O
{
for d in uniform32 uniform30 sorted; do
  python tools/sort_bench.py --pairs --dist $d --check
  python tools/sort_bench.py --dist $d --check
done
echo "--- COL_SORT_COLUMN=1"
for d in uniform32 sorted; do
  COL_SORT_COLUMN=1 python tools/sort_bench.py --pairs --dist $d
  COL_SORT_COLUMN=1 python tools/sort_bench.py --dist $d
done
python tools/sort_bench.py --pairs --n 1048576 --dist uniform30 --check
COL_SORT_COLUMN=1 python tools/sort_bench.py --pairs --n 1048576 --dist uniform30
} > $O/${TAG}_sort_ab.txt 2>&1
cat $O/${TAG}_sort_ab.txt
