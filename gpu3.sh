python -m pytest tests -m gpu -x -q 2>&1 | tail -5
SA_TIMING=1 python - <<'PY'
import time, sys
sys.path.insert(0,'.')
import signalalign_b200 as sa
from signalalign_b200 import synth
sa.gpu_init(0)
reads_d, table = synth.make_batch(2368, trim=0, block_mean=80)
model = sa.Model.from_tables("ACGT", 6, synth.nanopore_transitions(), table)
p = sa.default_params(threshold=0.01, expansion=100, trace_back=50, min_diags=1000)
reads = [sa.Read(r["target"], r["events"], r["anchors"], r["scale"], r["shift"], r["var"]) for r in reads_d]
for it in range(3):
    t=time.time(); b = sa.Batch(model, reads, p); t1=time.time(); b.run(); t2=time.time(); r=b.fetch(); t3=time.time(); b.close(); t4=time.time()
    print('prepare %.3f run %.3f fetch %.3f close %.3f' % (t1-t, t2-t1, t3-t2, t4-t3))
PY
nproc
