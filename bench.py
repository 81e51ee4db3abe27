#!/usr/bin/env python
"""bench.py -- the hot path of BASELINE.json on N B200s (or, with --impl reference, on the host CPU).

A "step" = one pass of the hot path over the synthetic structured-hex problem:
   fe_affine_operator%compute (batched integration + assembly into the device-built CSR) ->
   Jacobi set-up -> `--cg-iters` Jacobi-PCG iterations (SpMV + fused vector updates + reductions).
metric/value = CG SpMV bandwidth (algorithmic GB per SpMV / device time of the SpMV launches inside the
CG loop); the assembly throughput in MDOF/s of the same step is reported beside it (`assembly_mdof_per_s`).
`--kind` selects the integrand (BASELINE configs 2/5: poisson; config 4: vector_laplacian / elasticity).
Inputs (matrix 60-110 GB) are far larger than the 126 MB L2, so no explicit L2 flush is needed.
`--impl reference` times the reference's CPU algorithm (oracle port; the Fortran reference is unbuildable here) in its
MPI layout, one host thread per rank, on a bounded sample sized so that the requested steps finish in minutes.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

UNIT = "GB/s"
KINDS = {"poisson": 0, "vector_laplacian": 1, "elasticity": 2}
ELAST_MAT = [5.0 / 18.0, 5.0 / 12.0]       # lambda, mu (E = 1, nu = 0.2) as in the parity tests


def metric_name(order, kind):
    what = {"poisson": "Poisson", "vector_laplacian": "vector Laplacian", "elasticity": "linear elasticity"}[kind]
    return "cg_spmv_GBps (3D Q%d %s Jacobi-PCG; assembly MDOF/s reported beside it)" % (order, what)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--cells", type=int, default=0, help="cells per direction (0 = the BASELINE size of the kind/order if it fits)")
    ap.add_argument("--order", type=int, default=2)
    ap.add_argument("--kind", default="poisson", choices=sorted(KINDS))
    ap.add_argument("--cg-iters", type=int, default=25)
    ap.add_argument("--scatter", type=int, default=0)
    ap.add_argument("--cpu-cells", type=int, default=0, help="cells per direction of the bounded CPU sample (0 = sized by time)")
    ap.add_argument("--cpu-budget-s", type=float, default=0.0, help="CPU seconds the reference arm / cpu_baseline leg may spend in steps")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e-assembly", action="store_true")
    ap.add_argument("--subdomains", type=int, default=0,
                    help="number of FEMPAR sub-domains (parts); 0 = 8 (2x2x2, BASELINE config 3) when --gpus is 2, 4 or 8, else 1")
    ap.add_argument("--spmv-variant", type=int, default=0, help="0 default, 1 warp-per-row, 2 dictionary-compressed indices, 3 TMA tiles with ja (no pattern classes)")
    ap.add_argument("--asm-variant", type=int, default=0, help="0 default (row-owner gather where it applies + scatter), 1 generic kernel, 2 scatter kernel only")
    return ap.parse_args()


def spmv_bytes(n, nnz):
    # nnz*(8 val + 4 ja) + n*(8 irp + 8 y + 8 x)
    return nnz * 12 + n * 24


class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index, self.rows, self.stop = index, [], False
        self.t = threading.Thread(target=self.run, daemon=True)

    def run(self):
        while not self.stop:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                self.rows.append([c.strip() for c in out.strip().split(",")])
            except Exception:
                pass
            time.sleep(0.2)

    def __enter__(self):
        self.t.start()
        return self

    def __exit__(self, *a):
        self.stop = True
        self.t.join(timeout=6)

    def summary(self):
        sm = [float(r[0]) for r in self.rows if len(r) >= 7 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 7 and r[1].replace(".", "").isdigit()]
        reasons = set()
        for r in self.rows:
            if len(r) >= 7:
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ----------------------------------------------------------------------------------------------------------
# CPU side: the reference's algorithm (oracle port) on the host cores.  Only this part of bench.py touches oracle/.
# ----------------------------------------------------------------------------------------------------------
def host_info():
    model, l3 = "unknown", None
    try:
        for ln in open("/proc/cpuinfo"):
            if ln.startswith("model name"):
                model = ln.split(":", 1)[1].strip()
                break
    except Exception:
        pass
    try:
        seen = {}
        base = "/sys/devices/system/cpu"
        for cpu in os.listdir(base):
            p = os.path.join(base, cpu, "cache", "index3")
            if os.path.isdir(p):
                sz = open(os.path.join(p, "size")).read().strip()
                shared = open(os.path.join(p, "shared_cpu_list")).read().strip()
                mult = {"K": 1 << 10, "M": 1 << 20, "G": 1 << 30}.get(sz[-1], 1)
                seen[shared] = int(sz[:-1]) * mult if sz[-1] in "KMG" else int(sz)
        if seen:
            l3 = int(sum(seen.values()))
    except Exception:
        pass
    return model, l3


def usable_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def parts_for(cores):
    """largest power-of-two box split P <= cores (the partitions of par_test_poisson_fast: 2x2x2, 4x2x2, ...)."""
    p, dims, k = 1, [1, 1, 1], 0
    while p * 2 <= cores and p < 64:
        dims[k % 3] *= 2
        p *= 2
        k += 1
    return dims


def cpu_par_run(cells, order, kind, cg_iters, steps, warmup, dims):
    """P ranks as host threads, FEMPAR's par layout (oracle fo_par_run).  cells = per direction, a multiple of every dims[k]."""
    from concurrent.futures import ThreadPoolExecutor
    from oracle.oracle import OracleSpace, FUNC_LINEAR
    from oracle.par_oracle import par_run_threads
    ncomp = 1 if kind == "poisson" else 3
    P = int(np.prod(dims))
    t0 = time.perf_counter()

    def mk(p):
        s = OracleSpace([cells] * 3, order=order, ncomp=ncomp, parts=list(dims), part_id=(p + 1 if P > 1 else 0))
        s.build_pattern(False)
        return s
    with ThreadPoolExecutor(max_workers=P) as ex:          # ctypes releases the GIL: the parts are built concurrently
        spaces = list(ex.map(mk, range(P)))
    t_setup = time.perf_counter() - t0
    n = sum(s.num_free for s in spaces)
    nnz = sum(s.nnz for s in spaces)
    if P == 1:
        spaces[0].set_threads(1)
    r = par_run_threads(spaces, KINDS[kind], FUNC_LINEAR, None, ELAST_MAT if kind == "elasticity" else None, cg_iters, steps,
                        warmup)
    n_glob = ncomp * (order * cells - 1) ** 3
    out = dict(n=n, nnz=nnz, ranks=P, parts=list(dims), cells=cells, setup_s=t_setup,
               spmv_gbs=spmv_bytes(n, nnz) / r["apply_s"] / 1e9, asm_mdofs=n_glob / r["assembly_s"] / 1e6,
               ms_per_step=1e3 * (r["assembly_s"] + r["jacobi_s"] + r["pcg_s"]),
               checksum={"b_nrm2": r["b_nrm2"], "r_nrm2_after_iters": r["r_nrm2"], "diag_sum": r["diag_sum"]},
               matrix_bytes=nnz * 12 + n * 8)
    for s in spaces:
        s.close()
    return out


def cpu_pick_cells(order, kind, dims, step_budget_s, cores, cg_iters=25):
    """cells per direction of the bounded sample, sized by a calibration step on a small mesh (step time is linear in the
    number of DOFs) so that one step costs about `step_budget_s`; capped at 24 GB of matrix."""
    ncomp = 1 if kind == "poisson" else 3
    m = max(dims) * 2
    c = max(m, ((24 // order) // m) * m if ncomp == 1 else m)
    for _ in range(4):                                  # small meshes overstate the cost per cell: re-calibrate while growing
        r = cpu_par_run(c, order, kind, cg_iters, 1, 0, dims)
        per_cell_s = (r["ms_per_step"] * 1e-3) / c ** 3
        target = (step_budget_s / per_cell_s) ** (1.0 / 3.0)
        if target < 1.25 * c:
            c = int(target)
            break
        c = int(min(target, 2.0 * c))
        c = max(m, (c // m) * m)
    c = max(m, (c // m) * m)
    avg = {1: 27, 2: 64, 3: 125}.get(order, 216) * ncomp
    while c > m and ncomp * (order * c) ** 3 * avg * 12 > 24e9:
        c -= m
    return c


def cpu_sample_text(r, order, kind, cg_iters):
    return ("Q%d %s %d^3 cells (%d local DOFs over %d ranks [%s], %d nnz), %d PCG iterations/step, oracle C++ -O3, "
            "one host thread per rank" % (order, kind, r["cells"], r["n"], r["ranks"], "x".join(map(str, r["parts"])), r["nnz"],
                                          cg_iters))


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = usable_cores()
    dims = parts_for(cores)
    total = max(a.steps + a.warmup, 1)
    budget = a.cpu_budget_s if a.cpu_budget_s > 0 else 100.0
    cells = a.cpu_cells if a.cpu_cells else cpu_pick_cells(a.order, a.kind, dims, budget / total, cores, a.cg_iters)
    t0 = time.perf_counter()
    r = cpu_par_run(cells, a.order, a.kind, a.cg_iters, a.steps, a.warmup, dims)
    wall = time.perf_counter() - t0
    model, l3 = host_info()
    sample = cpu_sample_text(r, a.order, a.kind, a.cg_iters)
    line = {"impl": "reference", "metric": metric_name(a.order, a.kind), "value": r["spmv_gbs"], "unit": UNIT, "n_gpus": a.gpus,
            "steps": a.steps, "warmup": a.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "3D %s Q%d hex, Jacobi-PCG, CPU bounded sample of the GPU arm's workload: %s"
                                   % (a.kind, a.order, sample), "same_config": False, "cells": cells, "parts": r["parts"]},
            "assembly_mdof_per_s": r["asm_mdofs"], "checksum": r["checksum"],
            "cpu_baseline": {"value": r["spmv_gbs"], "unit": UNIT, "cores": r["ranks"], "kind": "port", "sample": sample,
                             "layout": "mpi (FEMPAR par_* layout, shared-memory mailboxes for the neighbour exchange)",
                             "host_cores_available": cores, "cpu_model": model, "l3_bytes": l3,
                             "matrix_bytes_over_l3": (r["matrix_bytes"] / l3) if l3 else None,
                             "setup_s": r["setup_s"], "wall_s": wall},
            "e2e": {"value": r["spmv_gbs"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


def cpu_baseline_leg(a):
    """rank 0, N = 1: ~10-30 s of CPU work.  Two variants: the serial path on one core and the MPI layout
    with one rank per host core."""
    cores = usable_cores()
    model, l3 = host_info()
    dims = parts_for(cores)
    budget = a.cpu_budget_s if a.cpu_budget_s > 0 else 12.0
    c_par = a.cpu_cells if a.cpu_cells else cpu_pick_cells(a.order, a.kind, dims, budget, cores, a.cg_iters)
    par = cpu_par_run(c_par, a.order, a.kind, a.cg_iters, 1, 0, dims)
    c_ser = cpu_pick_cells(a.order, a.kind, [1, 1, 1], budget * 0.5, 1, a.cg_iters)
    ser = cpu_par_run(c_ser, a.order, a.kind, a.cg_iters, 1, 0, [1, 1, 1])
    return {"value": par["spmv_gbs"], "unit": UNIT, "cores": par["ranks"], "kind": "port",
            "sample": cpu_sample_text(par, a.order, a.kind, a.cg_iters), "assembly_mdof_per_s": par["asm_mdofs"],
            "same_config": False, "cells": c_par, "cpu_model": model, "l3_bytes": l3, "host_cores_available": cores,
            "matrix_bytes_over_l3": (par["matrix_bytes"] / l3) if l3 else None,
            "mpi_layout": {"ranks": par["ranks"], "parts": par["parts"], "spmv_GBps": par["spmv_gbs"],
                           "assembly_mdof_per_s": par["asm_mdofs"], "ms_per_step": par["ms_per_step"], "checksum": par["checksum"]},
            "serial_1core": {"ranks": 1, "cells": c_ser, "spmv_GBps": ser["spmv_gbs"], "assembly_mdof_per_s": ser["asm_mdofs"],
                             "ms_per_step": ser["ms_per_step"], "sample": cpu_sample_text(ser, a.order, a.kind, a.cg_iters)}}


# ----------------------------------------------------------------------------------------------------------
# GPU arm
# ----------------------------------------------------------------------------------------------------------
def pick_cells(a, free_bytes, order, ncomp, nparts):
    if a.cells:
        return a.cells
    avg = {1: 27, 2: 64, 3: 125}.get(order, 216) * ncomp       # average nnz per row
    # BASELINE sizes: scalar Q2 256^3 (config 2), scalar Q3 128^3 (config 5), vector Q2 128^3 (config 4)
    first = 128 if (ncomp == 3 or order >= 3) else 256
    for c in [first] + [v for v in (224, 192, 160, 128, 96, 64, 32) if v < first]:
        n = ncomp * (order * c - 1) ** 3
        need = (n * avg * 12 + n * 8 * 12 + c ** 3 * ncomp * (order + 1) ** 3 * 4) / nparts
        if need * 1.08 < free_bytes:
            return c
    return 16


def asm_kernel_label(a, ncomp):
    if a.asm_variant == 1 or a.order > 3 or (a.kind == "elasticity" and a.order > 2):
        return "k_assemble_generic"
    sc = "k_assemble_fast<%d,%d,%s,%s>" % (a.order, ncomp, "atomic" if a.scatter == 0 else "colored",
                                           "elasticity" if a.kind == "elasticity" else "laplacian")
    if a.asm_variant == 0 and a.scatter == 0:
        return "k_assemble_rows<%d,%d> (row-owner gather, plain stores) + %s on the boundary rows" % (a.order, ncomp, sc)
    return sc


def main():
    a = parse()
    if a.impl == "reference":
        return run_reference(a)
    # Everything native libraries write to stdout during the run (e.g. "NCCL version ..." when NCCL_DEBUG=VERSION) goes to
    # stderr: stdout carries exactly ONE JSON line, printed by rank 0 at the end.
    sys.stdout.flush()
    saved_stdout = os.dup(1)
    os.dup2(2, 1)
    import torch
    import torch.distributed as dist
    import fempar_b200 as fb
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    fb.init(local)
    L = fb._lib.lib()
    fb._lib.check(L.fb2_set_spmv_variant(a.spmv_variant))
    fb._lib.check(L.fb2_set_assembly_variant(a.asm_variant))
    kind_id = KINDS[a.kind]
    ncomp = 1 if a.kind == "poisson" else 3
    mat = ELAST_MAT if a.kind == "elasticity" else None
    nsub = a.subdomains if a.subdomains else (8 if world > 1 else 1)
    nsub = max(nsub, world)
    parts = {1: (1, 1, 1), 2: (2, 1, 1), 4: (2, 2, 1), 8: (2, 2, 2)}[nsub]
    assert nsub % world == 0, "--subdomains must be a multiple of --gpus"
    ppr = nsub // world                     # co-resident sub-domains per process
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        fb.comm_init_from_torch(dist)
    fb.comm_set_parts_per_rank(ppr)
    free_b, total_b = torch.cuda.mem_get_info()
    cells = pick_cells(a, free_b, a.order, ncomp, world)
    N = [cells] * 3
    multi = ppr > 1
    t0 = time.perf_counter()
    if world == 1 and nsub == 1:
        my_parts = [0]
    else:
        my_parts = list(range(rank * ppr + 1, rank * ppr + ppr + 1))
    spaces = [fb.FeSpace(N, order=a.order, num_components=ncomp, num_parts_x_dim=parts, part_id=q) for q in my_parts]
    fb._lib.check(L.fb2_device_synchronize())
    t_space = time.perf_counter() - t0
    ops, t_pattern_ms = [], 0.0
    for sp in spaces:
        o = fb.FeAffineOperator()
        o.create(fb.csr_format, [False], [True], [1], sp, fb.DiscreteIntegration(kind_id, fb.func_linear, None, mat, a.scatter))
        t_pattern_ms += o.last_phase_ms("pattern")
        ops.append(o)
    space, op = spaces[0], ops[0]
    As = [o.get_matrix() for o in ops]
    A = As[0]
    n_loc, nnz_loc = sum(sp.num_free_dofs for sp in spaces), sum(m.nnz for m in As)
    ils = fb.IterativeLinearSolver(); ils.set_type_from_string("CG")
    ils.set_parameters_from_pl({fb.ILS_OUTPUT_FREQUENCY: 0})      # no convergence log inside the timed regions
    if multi:
        dinvs = [sp.create_dof_values() for sp in spaces]
        xs = [sp.create_dof_values() for sp in spaces]
    else:
        M = fb.JacobiPreconditioner(); M.create(op); M.symbolic_setup()
        ils.set_operators(A, M)
        x = space.create_dof_values()
        dinvs, xs = [M.inverse_diagonal], [x]

    def sync_all():
        fb._lib.check(L.fb2_device_synchronize())
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def step():
        if multi:
            t_asm = 0.0
            for o in ops:                                   # co-resident sub-domains are assembled one after the other
                o.force_compute()
                t_asm += o.last_phase_ms("integrate")
            bs = [o.get_translation() for o in ops]
            fb.vec_comm_multi(bs)                           # RHS comm()
            tj = time.perf_counter()
            fb.jacobi_setup_multi(As, dinvs)                # returns after a stream synchronize
            t_jac = (time.perf_counter() - tj) * 1e3
            t_tot, t_spmv = ils.iterate_multi(As, dinvs, bs, xs, a.cg_iters)
            return t_asm, t_tot, t_spmv, t_jac
        op.force_compute()                                  # re-assembly into the fixed pattern every step
        t_asm = op.last_phase_ms("integrate")
        tj = time.perf_counter()
        M.numerical_setup()                                 # returns after a stream synchronize
        t_jac = (time.perf_counter() - tj) * 1e3
        t_tot, t_spmv = ils.iterate(op.get_translation(), x, a.cg_iters)
        return t_asm, t_tot, t_spmv, t_jac

    sync_all()
    t0 = time.perf_counter()
    first = step()                                          # cold step: includes the one-off scatter/gather maps
    sync_all()
    t_first_ms = (time.perf_counter() - t0) * 1e3
    for _ in range(max(a.warmup - 1, 0)):
        step()
    sync_all()

    def phase(name):
        try:
            return op.last_phase_ms(name)
        except Exception:
            return None
    t_map_ms, t_gather_ms = phase("scatter_map"), phase("gather_plan")
    spmv_info = {"tiles": phase("spmv_tiles"), "tiles_in_pattern_class": phase("spmv_tiles_in_class"),
                 "tiles_private_descriptor": phase("spmv_tiles_private"), "transposed_rows": phase("spmv_transposed_rows"),
                 "pattern_classes": phase("spmv_classes"), "classes_setup_ms": phase("spmv_classes_ms")}
    asm_info = {"owner_cells_gather": phase("gather_fast_cells"), "cells_scatter": phase("gather_slow_cells"),
                "rows_scatter": phase("gather_slow_rows")}
    # per-rank timeline of the overlapped par apply (FB2_PAR_TIMELINE=1): gathered from every rank after the timed steps
    asm_ms, cg_ms, spmv_ms, jac_ms = [], [], [], []
    with ClockSampler(local) as clk:
        t0 = time.perf_counter()
        for _ in range(a.steps):
            ta, tt, ts, tj = step()
            asm_ms.append(ta); cg_ms.append(tt); spmv_ms.append(ts); jac_ms.append(tj)
        sync_all()
        wall = time.perf_counter() - t0
    # device-time per step (events on the library stream), max over ranks
    dev = torch.tensor([np.mean(asm_ms), np.mean(cg_ms), np.mean(spmv_ms) / a.cg_iters, wall / a.steps * 1e3, np.mean(jac_ms),
                        t_first_ms, t_space * 1e3, t_pattern_ms],
                       dtype=torch.float64, device="cuda")
    tot = torch.tensor([float(n_loc), float(nnz_loc)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(dev, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    asm_t, cg_t, spmv_t, wall_t, jac_t, first_t, space_t, pattern_t = [float(v) for v in dev.cpu()]
    n_glob_rows, nnz_glob = [float(v) for v in tot.cpu()]
    n_free_global = ncomp * (a.order * cells - 1) ** 3
    bytes_all = spmv_bytes(n_glob_rows, nnz_glob)
    value = bytes_all / (spmv_t * 1e-3) / 1e9
    asm_mdofs = n_free_global / (asm_t * 1e-3) / 1e6

    # checksums of the state the timed steps left behind (same 8-part problem at N = 2, 4, 8; same global problem at N = 1):
    # |b|_2, |b - A x|_2 after the PCG iterations of the last step, sum of the assembled (comm'ed) diagonal
    bs = [o.get_translation() for o in ops]
    tmp = [sp.create_dof_values() for sp in spaces]
    ones = [sp.create_dof_values() for sp in spaces]
    if multi:
        chk_b = np.sqrt(fb.vec_dot_multi(bs, bs))
        fb.spmv_multi(As, xs, tmp)
        for t_, b_ in zip(tmp, bs):
            t_.axpby(-1.0, b_, 1.0)
        chk_r = np.sqrt(fb.vec_dot_multi(tmp, tmp))
        for t_, d_, o_ in zip(tmp, dinvs, ones):
            t_.copy(d_); t_.entrywise_invert(); o_.init(1.0)
        chk_d = fb.vec_dot_multi(tmp, ones)
    else:
        chk_b = bs[0].nrm2()
        A.apply(xs[0], tmp[0])
        tmp[0].axpby(-1.0, bs[0], 1.0)
        chk_r = tmp[0].nrm2()
        tmp[0].copy(dinvs[0]); tmp[0].entrywise_invert(); ones[0].init(1.0)
        chk_d = tmp[0].dot(ones[0])
    for v in tmp + ones:
        v.free()
    checksum = {"b_nrm2": chk_b, "r_nrm2_after_iters": chk_r, "diag_sum": chk_d}

    # e2e: the same PCG iterations through the C-ABI call with HOST buffers (H2D of b, D2H of x inside the timed region)
    ils.set_parameters_from_pl({fb.ILS_RTOL: 0.0, fb.ILS_ATOL: -1.0, fb.ILS_MAX_NUM_ITERATIONS: a.cg_iters})
    ils.check_every = a.cg_iters
    ts = []
    if multi:
        b_hosts = [torch.from_numpy(o.get_translation().get_entries()).pin_memory() for o in ops]
        x_hosts = [torch.empty(sp.num_free_dofs, dtype=torch.float64).pin_memory() for sp in spaces]
        bvs = [sp.create_dof_values() for sp in spaces]
        for k in range(3):
            sync_all()
            t0 = time.perf_counter()
            for bv, bh in zip(bvs, b_hosts):
                bv.set_entries(bh.numpy())                  # H2D
            ils.apply_multi(As, dinvs, bvs, xs)
            for xv, xh in zip(xs, x_hosts):
                fb._lib.check(L.fb2_vec_get(xv.h, xh.data_ptr()))   # D2H
            ts.append(time.perf_counter() - t0)
    else:
        b_host = torch.empty(n_loc, dtype=torch.float64).pin_memory()
        b_host.copy_(torch.from_numpy(op.get_translation().get_entries()))
        x_host = torch.empty(n_loc, dtype=torch.float64).pin_memory()
        for k in range(3):
            sync_all()
            t0 = time.perf_counter()
            ils.apply(b_host.numpy(), x_host.numpy())
            fb._lib.check(L.fb2_device_synchronize())
            ts.append(time.perf_counter() - t0)
    te = torch.tensor([min(ts[1:])], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e = {"value": a.cg_iters * bytes_all / float(te.item()) / 1e9, "unit": UNIT, "h2d_bytes_per_step": int(n_glob_rows * 8),
           "d2h_bytes_per_step": int(n_glob_rows * 8),
           "what": ("fb2_pcg_solve_host" if not multi else "fb2_vec_set + fb2_pcg_solve_multi + fb2_vec_get") +
                   " on every rank: H2D b + %d PCG iterations + D2H x (max over ranks)" % a.cg_iters}
    # e2e of fe_affine_operator%compute with HOST buffers: the user's source-term samples at the quadrature points go H2D
    # (fb2_fe_space_set_source_values), the operator is integrated and assembled, the translation (RHS) comes back D2H
    e2e_asm = None
    if a.kind == "poisson" and not a.no_e2e_assembly:
        nq = (a.order + 1) ** 3
        srcs = [torch.zeros(sp.num_local_cells * nq, dtype=torch.float64).pin_memory() for sp in spaces]
        rhs_h = [torch.empty(sp.num_free_dofs, dtype=torch.float64).pin_memory() for sp in spaces]
        tsa = []
        for k in range(3):
            sync_all()
            t0 = time.perf_counter()
            for sp, o, sh, rh in zip(spaces, ops, srcs, rhs_h):
                fb._lib.check(L.fb2_fe_space_set_source_values(sp.h, sh.data_ptr()))          # H2D
                o.force_compute()
                fb._lib.check(L.fb2_vec_get(o.get_translation().h, rh.data_ptr()))            # D2H
            tsa.append(time.perf_counter() - t0)
        for sp in spaces:
            sp.set_source_values(None)
        ta_ = torch.tensor([min(tsa[1:])], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(ta_, op=dist.ReduceOp.MAX)
        e2e_asm = {"assembly_mdof_per_s": n_free_global / float(ta_.item()) / 1e6, "ms": float(ta_.item()) * 1e3,
                   "h2d_bytes_per_step": int(sum(s_.numel() for s_ in srcs) * 8 * world),
                   "d2h_bytes_per_step": int(n_glob_rows * 8),
                   "what": "fb2_fe_space_set_source_values (H2D f(x_q) samples) + fb2_integrate_assemble + fb2_vec_get of the RHS"}

    timeline = None
    if world > 1 and os.environ.get("FB2_PAR_TIMELINE", "0") != "0":
        names = ("par_itf_ms", "par_exchange_ms", "par_interior_ms", "par_exposed_wait_ms", "par_rest_ms")
        mine = torch.tensor([phase(k) or 0.0 for k in names] + [float(np.mean(spmv_ms)) / a.cg_iters, float(np.mean(cg_ms)) / a.cg_iters],
                            dtype=torch.float64, device="cuda")
        allr = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(allr, mine)
        timeline = {"columns": list(names) + ["spmv_apply_ms", "pcg_iteration_ms"], "per_rank": [[float(v) for v in t.cpu()] for t in allr]}
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    # DRAM bytes per SpMV launch from the committed ncu capture of this workload (profiles/spmv_traffic.json), else null
    traffic = None
    try:
        tr = json.load(open(os.path.join(ROOT, "profiles", "spmv_traffic.json")))
        key = "%s_Q%d_%d_parts%d_v%d" % (a.kind, a.order, cells, world, a.spmv_variant)
        if key in tr:
            traffic = tr[key]["dram_bytes_per_launch"]
    except Exception:
        pass
    per_gpu_bytes = spmv_bytes(n_loc, nnz_loc)
    class_share = 0.0
    if a.spmv_variant == 0 and spmv_info["tiles"] and spmv_info["tiles_in_pattern_class"] is not None and ppr == 1:
        class_share = spmv_info["tiles_in_pattern_class"] / spmv_info["tiles"]
        # tiles with a private descriptor stream a uint16 index instead of ja: 10 instead of 12 B per non-zero
        class_share += 0.5 * (spmv_info.get("tiles_private_descriptor") or 0.0) / spmv_info["tiles"]
    achieved = per_gpu_bytes / (np.mean(spmv_ms) / a.cg_iters * 1e-3) / 1e9
    asm_bytes = nnz_loc * 8 + n_loc * 8
    try:
        spmv_label = {0: "k_spmv_stream (TMA-streamed CSR row tiles; column indices of verified pattern classes rebuilt from class "
                         "tables, ja streamed for the rest; fused <Ap,p>)", 1: "k_spmv<true> (warp-per-row)",
                      2: "k_spmv_stream, dictionary-compressed indices", 3: "k_spmv_stream with ja streamed for every tile"}[a.spmv_variant]
    except KeyError:
        spmv_label = "k_spmv_stream variant %d" % a.spmv_variant
    ms_step = asm_t + jac_t + cg_t
    line = {"metric": metric_name(a.order, a.kind), "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": "3D %s Q%d hex %d^3 cells (%d free DOFs, %d nnz), u=x+y+z Dirichlet, Jacobi-PCG; "
                                   "step = integrate+assemble + Jacobi setup + %d PCG iterations"
                                   % (a.kind, a.order, cells, n_free_global, int(nnz_glob), a.cg_iters),
                       "kind": a.kind, "order": a.order, "cells": cells,
                       "parts": list(parts), "subdomains_per_gpu": ppr, "l2": "inputs larger than L2 (no flush needed)",
                       "scatter": a.scatter, "asm_variant": a.asm_variant, "spmv_variant": a.spmv_variant},
            "assembly_mdof_per_s": asm_mdofs, "assembly_ms": asm_t, "jacobi_setup_ms": jac_t,
            "pcg_ms_per_iteration": cg_t / a.cg_iters, "spmv_ms": spmv_t,
            "ms_per_step_parts": {"assembly": asm_t, "jacobi_setup": jac_t, "pcg": cg_t},
            "setup_ms": {"fe_space": space_t, "csr_pattern": pattern_t, "scatter_map": t_map_ms, "gather_plan": t_gather_ms,
                         "spmv_pattern_classes": spmv_info["classes_setup_ms"]},
            "spmv_tiles": spmv_info, "assembly_cells": asm_info,
            "peer_allreduce": bool(phase("peer_allreduce")) if world > 1 else None,
            # one-shot create -> compute -> solve (what every reference driver does): set-up + the first (cold) step
            "cold_start_ms": space_t + pattern_t + first_t,
            "wall_ms_per_step": wall_t, "checksum": checksum,
            "roofline": {"bound": "hbm", "kernel": spmv_label, "achieved": achieved,
                         "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                         "algorithmic_bytes_per_launch": per_gpu_bytes,
                         "frac_of_nominal_8000_GBps": achieved / 8000.0,
                         # tiles of a verified pattern class stream no column indices: what the kernel actually has to move
                         "streamed_bytes_per_launch_estimate": per_gpu_bytes - 4.0 * nnz_loc * class_share,
                         "streamed_GBps": (per_gpu_bytes - 4.0 * nnz_loc * class_share) / (np.mean(spmv_ms) / a.cg_iters * 1e-3) / 1e9,
                         # DRAM bytes of the ncu capture over this run's launch time: the physical bandwidth the kernel draws
                         "traffic_GBps": (traffic / (np.mean(spmv_ms) / a.cg_iters * 1e-3) / 1e9) if traffic else None,
                         "traffic_frac_of_peak": (traffic / (np.mean(spmv_ms) / a.cg_iters * 1e-3) / 1e9 / peak) if traffic else None,
                         "peak_source": "measured (MEASURED_PEAKS.json hbm_gbs)" if peaks else "fallback 6650"},
            "roofline_assembly": {"bound": "hbm", "kernel": asm_kernel_label(a, ncomp),
                                  "achieved": asm_bytes / (np.mean(asm_ms) * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                                  "frac": asm_bytes / (np.mean(asm_ms) * 1e-3) / 1e9 / peak,
                                  "algorithmic_bytes_per_launch": asm_bytes},
            # BLAS-1 traffic of one PCG iteration: the reference's unfused vector TBPs move 18 n doubles, the fused
            # kernels here 11 n (x/r update with four partial dots: 4 reads + 2 writes + dinv; p update: 3 reads + 1 write)
            "pcg_vector_bytes_per_iteration": {"reference_unfused": int(18 * 8 * n_glob_rows), "fused": int(11 * 8 * n_glob_rows),
                                               "measured_ms": (cg_t / a.cg_iters) - spmv_t},
            "clocks": clk.summary(), "e2e": e2e, "e2e_assembly": e2e_asm, "par_timeline": timeline,
            # kernels of libfempar_b200.so per step: Dirichlet interpolation + memsets/assembly (4), Jacobi set-up (2), PCG set-up
            # (9), then per PCG iteration: serial 6 (SpMV, 2 partial reductions, x/r update, scalars, p update); par 14 (two
            # SpMV launches, 4 pack/unpack kernels, interface dot, reductions, updates; NCCL's own kernels not counted)
            "gpu_launches": a.steps * ((4 + 2) * ppr + 9 + (6 if (world == 1 and not multi) else 14 * ppr) * a.cg_iters)}
    if not a.no_cpu_baseline and world == 1:       # reported on rank 0 at N=1 only
        line["cpu_baseline"] = cpu_baseline_leg(a)
    sys.stdout.flush()
    os.dup2(saved_stdout, 1)
    print(json.dumps(line), flush=True)
    os.dup2(2, 1)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
