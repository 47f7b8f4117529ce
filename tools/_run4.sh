timeout 900 python -m pytest tests -q -m gpu -k "edit or c4 or golden or oversized" -x 2>&1 | tail -3
timeout 300 python tools/stage_table.py 3 20000000 3 2>&1 | grep -E "edit|sum|c4|filter|table_ranks"
