"""Bit-reproducibility probe: SHA-256 of the C3 delta of one damped trial, printed per process (run it several times)."""
import hashlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from gtsam_petercdev_b200.engine import Engine
w = sys.argv[1] if len(sys.argv) > 1 else "c3"
p = bench.make_problem(w)
e = Engine(p, device=0)
e.set_ordering(bench.ordering_for(p, w, "metis"))
out = []
for _ in range(2):
    e.linearize(); e.factorize(1e-5); e.backsolve()
    out.append(hashlib.sha256(e.delta().tobytes()).hexdigest()[:16])
print(w, out)
