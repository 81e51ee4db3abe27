"""GPU parity of the par_* layer: every part of a partition is built on the device (one at a time on cuda:0) and compared
bit-exactly with the oracle's restatement of FEMPAR's par_fe_space_t / dof_import_t; the NCCL path (halo exchange,
weighted dots, par PCG) runs under torchrun when the box has >= 2 GPUs."""
import os
import subprocess
import sys

import numpy as np
import pytest

from oracle.oracle import OracleSpace, KIND_POISSON, FUNC_LINEAR

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

PARTITIONS = [((4, 4, 4), 1, 1, (2, 1, 1)), ((8, 8, 8), 1, 1, (2, 2, 1)), ((6, 6, 6), 2, 1, (2, 2, 2)), ((6, 6, 6), 1, 1, (3, 3, 3)),
              ((6, 4, 6), 3, 1, (2, 1, 3)), ((4, 4, 4), 2, 3, (2, 2, 1)), ((8, 4, 4), 2, 1, (4, 2, 1))]


@pytest.mark.parametrize("N,p,ncomp,parts", PARTITIONS)
def test_par_layer_bit_exact(fb2, N, p, ncomp, parts):
    for part in range(1, int(np.prod(parts)) + 1):
        s = fb2.FeSpace(N, order=p, num_components=ncomp, num_parts_x_dim=parts, part_id=part)
        o = OracleSpace(N, order=p, ncomp=ncomp, parts=list(parts), part_id=part)
        assert (s.num_cells, s.num_local_cells, s.num_free_dofs, s.num_free_ghost_dofs, s.num_dirichlet_dofs) == \
               (o.num_cells, o.num_local_cells, o.num_free, o.num_free_ghost, o.num_dirichlet)
        g, pt = s.get_cells_gids()
        assert np.array_equal(g, o.cells_gid()) and np.array_equal(pt, o.cells_part())
        assert (s.num_interior_dofs, s.num_interface_dofs) == (o.num_interior, o.num_interface)
        assert np.array_equal(s.get_lst_dofs_gids(), o.lst_dofs_gids())          # after the I-then-G renumbering
        d, od = s.get_dof_import(), o.dof_import()
        for k in ("list_rcv", "list_snd", "unpack_idx", "pack_idx"):
            assert np.array_equal(d[k], od[k]), (part, k)
        assert np.array_equal(d["rcv_ptrs"] - 1, od["rcv_ptrs"]) and np.array_equal(d["snd_ptrs"] - 1, od["snd_ptrs"])
        assert np.array_equal(d["weight"], od["weight"])
        s.free()


@pytest.mark.parametrize("N,p,parts", [((6, 6, 6), 2, (2, 2, 2)), ((8, 8, 8), 1, (2, 2, 1))])
def test_par_local_matrix_matches_oracle(fb2, N, p, parts):
    """sub-assembled (Neumann) local matrices and local RHS before comm(): graph bit-exact, values to 1e-12."""
    for part in range(1, int(np.prod(parts)) + 1):
        s = fb2.FeSpace(N, order=p, num_parts_x_dim=parts, part_id=part)
        op = fb2.FeAffineOperator()
        op.create(fb2.csr_format, [False], [True], [1], s, fb2.DiscreteIntegration(fb2.poisson, fb2.func_linear))
        op.compute()
        o = OracleSpace(N, order=p, parts=list(parts), part_id=part)
        o.assemble(KIND_POISSON, FUNC_LINEAR, route=1)
        irp, ja = op.get_matrix().get_pattern()
        assert np.array_equal(irp - 1, o.irp()) and np.array_equal(ja, o.ja())
        v, ov = op.get_matrix().get_values(), o.val()
        assert np.max(np.abs(v - ov)) <= 1e-12 * np.max(np.abs(ov))
        b, ob = op.get_translation().get_entries(), o.rhs()     # no communicator: the local (unassembled) RHS
        assert np.max(np.abs(b - ob)) <= 1e-12 * max(np.max(np.abs(ob)), 1e-300)


def _ngpus():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


@pytest.mark.parametrize("world", [2, 4, 8])
def test_par_pcg_nccl(world):
    if _ngpus() < world:
        pytest.skip("needs %d GPUs" % world)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world), "--master-addr", "127.0.0.1",
           "--master-port", str(29500 + world), os.path.join(ROOT, "tests", "par_nccl_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "PAR_NCCL_OK" in r.stdout


@pytest.mark.parametrize("world", [1, 2, 4])
def test_par_8_parts_on_fewer_gpus_nccl(world):
    """BASELINE config 3: the 8-sub-domain layout on 1, 2 and 4 GPUs (8 / 4 / 2 co-resident parts per process)."""
    if _ngpus() < world:
        pytest.skip("needs %d GPUs" % world)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world), "--master-addr", "127.0.0.1",
           "--master-port", str(29600 + world), os.path.join(ROOT, "tests", "par_multi_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "PAR_MULTI_OK" in r.stdout
