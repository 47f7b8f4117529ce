for v in 0 1 2 3 4; do
  echo "variant $v"
  HUMID_GPU_RS_VARIANT=$v timeout 120 python tools/stage_table.py 1 50000000 4 2>&1 | grep -E "table_partition|sort_words|join_sort|^sum|clusters"
done
