"""Assembly time of a host-numbered (uploaded) Q2 space vs the device-numbered structured space on the same mesh."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import fempar_b200 as fb2
from oracle import hex_oracle as ho

fb2.init(0)
for p, N in ((1, 128), (2, 96), (3, 48)):
    s0 = fb2.FeSpace([N] * 3, order=p)
    gids = s0.get_lst_dofs_gids()
    X = ho.structured_cell_coords(N, lambda X: X + 0.01 * np.sin(np.pi * X) * np.sin(np.pi * X[..., [1, 2, 0]]))
    s1 = fb2.HostNumberedFeSpace(p, 1, gids, X, s0.num_free_dofs, s0.num_dirichlet_dofs)
    for name, s in (("structured", s0), ("host-numbered", s1)):
        for variant in (0, 1):
            fb2._lib.check(fb2._lib.lib().fb2_set_assembly_variant(variant))
            op = fb2.FeAffineOperator()
            t0 = time.time()
            op.create(fb2.csr_format, [False], [True], [1], s, fb2.DiscreteIntegration(fb2.poisson, fb2.func_linear, None, None, fb2.scatter_atomic))
            pat = op.last_phase_ms("pattern")
            op.compute()
            ms = []
            for _ in range(3):
                op.force_compute()
                ms.append(op.last_phase_ms("integrate"))
            print("Q%d %d^3 %-14s variant %d: pattern %.1f ms, scatter map %.1f ms, integrate+assemble %.2f ms (%.1f MDOF/s), nnz %d" % (
                p, N, name, variant, pat, op.last_phase_ms("scatter_map") if variant == 0 else 0.0, min(ms),
                s.num_free_dofs / min(ms) / 1e3, op.get_matrix().nnz), flush=True)
            op.free()
