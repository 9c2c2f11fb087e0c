python - <<'PY'
import sys, os
sys.path.insert(0,'tools'); import synthetic as syn
p = syn.write_scene('/tmp/syn_ab', 5, min_dist=0.0035, film=64, spp=1)
PY
python tools/mo_windows.py /tmp/syn_ab/ico5.scene 2>&1 | tail -2
