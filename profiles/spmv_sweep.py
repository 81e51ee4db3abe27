"""SpMV tile-geometry sweep on one matrix (profiles/): for every (R, STAGES, CAP, CHUNK) the tile table and pattern classes are
rebuilt (fb2_csr_spmv_reset) and the SpMV launches inside 12 PCG iterations are timed with CUDA events (fb2_pcg_iterate).
usage: python profiles/spmv_sweep.py [cells] [order]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import fempar_b200 as fb  # noqa: E402

cells = int(sys.argv[1]) if len(sys.argv) > 1 else 256
order = int(sys.argv[2]) if len(sys.argv) > 2 else 2
fb.init(0)
L = fb._lib.lib()
s = fb.FeSpace([cells] * 3, order=order)
op = fb.FeAffineOperator()
op.create(fb.csr_format, [False], [True], [1], s, fb.DiscreteIntegration(fb.poisson, fb.func_linear))
op.compute()
A = op.get_matrix()
M = fb.JacobiPreconditioner(); M.create(op); M.numerical_setup()
ils = fb.IterativeLinearSolver(); ils.set_type_from_string("CG"); ils.set_operators(A, M)
x = s.create_dof_values()
bytes_alg = A.nnz * 12 + A.num_rows * 24
# (R, STAGES, CAP, CHUNK, CONSUMERS, TRANSPOSE); round-2 call 4 showed 2 stages x 4 CTAs/SM ahead of deeper rings and smaller tiles;
# TRANSPOSE: row map of the consumer lane groups (0 natural, 1 transposed + staggered chunks), A/B/A/B against box drift
# WAIT_NS: suspend-time hint of the consumers' mbarrier wait (0 = plain try_wait spin)
# last columns: FB2_SPMV_L2HINT (bit 0: val / ja copies evict_first, bit 1: x segments evict_last), FB2_SPMV_XALIGN (x-window
# segments placed so that window position == column modulo 16 doubles)
combos = [(32, 2, 2304, 32, 256, 1, 0, 1, 1), (32, 2, 2304, 32, 256, 1, 0, 1, 0), (32, 2, 2304, 32, 256, 1, 0, 1, 1), (32, 2, 2304, 32, 256, 1, 0, 1, 0),
          (32, 2, 2304, 32, 256, 0, 0, 1, 1), (32, 2, 2304, 32, 256, 0, 0, 1, 0)]
if order != 2:
    combos = [(16, 2, 2304, 32, 128, 0, 0, 1, 1), (16, 2, 2304, 32, 128, 0, 0, 1, 0), (16, 2, 2304, 32, 128, 0, 0, 1, 1), (16, 2, 2304, 32, 128, 0, 0, 1, 0)]
for R, st, cap, chunk, nc, tr, wait_ns, l2hint, xalign in combos:
    os.environ["FB2_SPMV_L2HINT"] = str(l2hint)
    os.environ["FB2_SPMV_XALIGN"] = str(xalign)
    os.environ["FB2_SPMV_R"], os.environ["FB2_SPMV_STAGES"], os.environ["FB2_SPMV_CAP"], os.environ["FB2_SPMV_CHUNK"] = \
        str(R), str(st), str(cap), str(chunk)
    os.environ["FB2_SPMV_CONSUMERS"] = str(nc)
    os.environ["FB2_SPMV_TRANSPOSE"] = str(tr)
    os.environ["FB2_SPMV_WAIT_NS"] = str(wait_ns)
    try:
        fb._lib.check(L.fb2_csr_spmv_reset(A.h))
        ils.iterate(op.get_translation(), x, 3)
        t, ts = ils.iterate(op.get_translation(), x, 12)
        ms = ts / 12
        print("transpose=%d l2hint=%d xalign=%d nc=%d " % (tr, l2hint, xalign, nc), end="")
        print("R=%-3d stages=%d cap=%-5d chunk=%-4d  spmv %.3f ms  %.0f GB/s algorithmic  tiles_in_class %.0f / %.0f  ctas/sm %.0f  "
              "row-map cost natural %.4g transposed %.4g" % (
                  R, st, cap, chunk, ms, bytes_alg / ms / 1e6, op.last_phase_ms("spmv_tiles_in_class"), op.last_phase_ms("spmv_tiles"),
                  op.last_phase_ms("spmv_ctas_per_sm"), op.last_phase_ms("spmv_row_map_cost_natural"),
                  op.last_phase_ms("spmv_row_map_cost_transposed")), flush=True)
    except Exception as e:   # a geometry that does not fit is reported, not fatal
        print("R=%d stages=%d cap=%d chunk=%d transpose=%d: %s" % (R, st, cap, chunk, tr, e), flush=True)
