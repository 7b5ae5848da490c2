for d in 0; do echo "== debug $d"; TSM_TC_DEBUG=$d PYTHONPATH=. python tools/mb_sweep.py tc 2>&1 | grep -v "tiles/cta=16\|forward raw"; done
