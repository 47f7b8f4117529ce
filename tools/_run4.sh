timeout 900 python -m pytest tests -q -m gpu -k "edit or c4 or golden" -x 2>&1 | tail -5
timeout 300 python tools/stage_table.py 3 20000000 3 2>&1 | grep -E "edit|sum|c4"
