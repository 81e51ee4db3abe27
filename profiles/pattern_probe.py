"""CSR graph set-up time (count + scan + fill) of the lattice builder; FB2_ROW_FILL=0 selects the shared-memory row sort."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import fempar_b200 as fb2
fb2.init(0)
for p, N in ((2, 64), (2, 128), (2, 256), (1, 256), (3, 96)):
    s = fb2.FeSpace([N] * 3, order=p)
    for rep in range(2):
        A = fb2.SparseMatrix(s)
        ms = fb2._lib.C.c_double(0)
        A.L.fb2_last_phase_ms(b"pattern", fb2._lib.C.byref(ms))
        print("Q%d %d^3 rep %d: pattern %.1f ms, nnz %d" % (p, N, rep, ms.value, A.nnz), flush=True)
        A.free()
    s.free()
