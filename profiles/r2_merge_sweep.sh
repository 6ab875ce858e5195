timeout 150 python -m pytest tests/test_gpu_merge.py -m gpu -q -x 2>&1 | tail -3
for g in 2 4 8; do for e in 1 0; do echo "G=$g early=$e"; FM_MERGE_EARLY_ROWS=$e timeout 100 python profiles/merge_virtual_bench.py $g 5; done; done
for b in 4 16; do echo "blocks/SM=$b"; FM_MERGE_BLOCKS_PER_SM=$b timeout 100 python profiles/merge_virtual_bench.py 8 5; done
