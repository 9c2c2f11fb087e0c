python - <<'PY'
import sys, os
sys.path.insert(0,'tools'); import synthetic as syn
p = syn.write_scene('/tmp/syn_ab', 8, min_dist=0.0035, film=2048, spp=1)
print(p)
PY
MO_AB_VARIANTS=0 python tools/mo_table_ab.py /tmp/syn_ab/ico8.scene > gpurun_out/ab_syn.log 2>&1; tail -2 gpurun_out/ab_syn.log
