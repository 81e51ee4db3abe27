#!/usr/bin/env python
"""bench.py -- the hot path of BASELINE.json on N B200s (or, with --impl reference, on the host CPU).

A "step" = one pass of the hot path over the synthetic structured-hex problem:
   fe_affine_operator%compute (batched integration + scatter into the device-built CSR) ->
   Jacobi set-up -> `--cg-iters` Jacobi-PCG iterations (SpMV + fused vector updates + reductions).
metric/value = CG SpMV bandwidth (algorithmic GB per SpMV / device time of the SpMV launches inside the
CG loop); the assembly throughput in MDOF/s of the same step is reported beside it (`assembly_mdof_per_s`).
Inputs (matrix 100+ GB) are far larger than the 126 MB L2, so no explicit L2 flush is needed.
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

METRIC = "cg_spmv_GBps (3D Q2 Poisson Jacobi-PCG; assembly MDOF/s reported beside it)"
UNIT = "GB/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--cells", type=int, default=0, help="cells per direction (0 = 256 if it fits, else largest that fits)")
    ap.add_argument("--order", type=int, default=2)
    ap.add_argument("--cg-iters", type=int, default=25)
    ap.add_argument("--scatter", type=int, default=0)
    ap.add_argument("--cpu-cells", type=int, default=48, help="cells per direction of the bounded CPU sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--subdomains", type=int, default=0,
                    help="number of FEMPAR sub-domains (parts); 0 = 8 (2x2x2, BASELINE config 3) when --gpus is 2, 4 or 8, else 1")
    ap.add_argument("--spmv-variant", type=int, default=0, help="0 TMA-streamed tiles, 1 warp-per-row, 2 streamed tiles with dictionary-compressed column indices")
    ap.add_argument("--asm-variant", type=int, default=0, help="0 register-blocked contraction + scatter map, 1 generic kernel")
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


def cpu_reference_run(cells, order, cg_iters, steps, warmup):
    """The reference's CPU path (oracle port; the Fortran reference cannot be built here) on a bounded sample."""
    from oracle.oracle import OracleSpace, KIND_POISSON, FUNC_LINEAR
    t0 = time.perf_counter()
    s = OracleSpace([cells] * 3, order=order)
    cores = s.set_threads(0)       # SpMV rows split over all host threads; the cell loop is serial like the reference's
    t_space = time.perf_counter() - t0
    s.build_pattern(False)
    n, nnz = s.num_free, s.nnz
    asm, spm, stp = [], [], []
    for it in range(warmup + steps):
        t0 = time.perf_counter()
        s.assemble(KIND_POISSON, FUNC_LINEAR, route=1)
        t1 = time.perf_counter()
        s.jacobi_setup()
        x = np.random.default_rng(0).standard_normal(n)
        t2 = time.perf_counter()
        for _ in range(cg_iters):
            x = s.spmv(x)
        t3 = time.perf_counter()
        r = s.cg(prec=1, rtol=0.0, atol=-1.0, maxit=cg_iters)
        t4 = time.perf_counter()
        if it >= warmup:
            asm.append(t1 - t0); spm.append((t3 - t2) / cg_iters); stp.append((t1 - t0) + (t4 - t3))
    gbs = spmv_bytes(n, nnz) / np.mean(spm) / 1e9
    return dict(n=n, nnz=nnz, cores=cores, spmv_gbs=gbs, asm_mdofs=n / np.mean(asm) / 1e6, ms_per_step=1e3 * np.mean(stp),
                sample="Q%d %d^3 cells (%d DOFs, %d nnz), %d PCG iterations/step, oracle C++ -O3" % (order, cells, n, nnz, cg_iters))


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    r = cpu_reference_run(a.cpu_cells, a.order, a.cg_iters, max(1, min(a.steps, 3)), min(a.warmup, 1))
    line = {"impl": "reference", "metric": METRIC, "value": r["spmv_gbs"], "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "3D Poisson Q%d hex, Jacobi-PCG, CPU bounded sample: %s" % (a.order, r["sample"])},
            "assembly_mdof_per_s": r["asm_mdofs"],
            "cpu_baseline": {"value": r["spmv_gbs"], "unit": UNIT, "cores": r["cores"], "kind": "port", "sample": r["sample"]},
            "e2e": {"value": r["spmv_gbs"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


def pick_cells(a, free_bytes, order, nparts):
    if a.cells:
        return a.cells
    avg = {1: 27, 2: 64, 3: 125}.get(order, 216)       # average nnz per row
    for c in (256, 224, 192, 160, 128, 96, 64, 32):
        n = (order * c - 1) ** 3
        need = (n * avg * 12 + n * 8 * 12 + c ** 3 * (order + 1) ** 3 * 4) / nparts
        if need * 1.08 < free_bytes:
            return c
    return 16


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
    fb._lib.check(fb._lib.lib().fb2_set_spmv_variant(a.spmv_variant))
    fb._lib.check(fb._lib.lib().fb2_set_assembly_variant(a.asm_variant))
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
    cells = pick_cells(a, free_b, a.order, world)
    N = [cells] * 3
    multi = ppr > 1
    t0 = time.perf_counter()
    if world == 1 and nsub == 1:
        my_parts = [0]
    else:
        my_parts = list(range(rank * ppr + 1, rank * ppr + ppr + 1))
    spaces = [fb.FeSpace(N, order=a.order, num_parts_x_dim=parts, part_id=q) for q in my_parts]
    t_space = time.perf_counter() - t0
    ops, t_pattern_ms = [], 0.0
    for sp in spaces:
        o = fb.FeAffineOperator()
        o.create(fb.csr_format, [False], [True], [1], sp, fb.DiscreteIntegration(fb.poisson, fb.func_linear, None, None, a.scatter))
        t_pattern_ms += o.last_phase_ms("pattern")
        ops.append(o)
    space, op = spaces[0], ops[0]
    As = [o.get_matrix() for o in ops]
    A = As[0]
    n_loc, nnz_loc = sum(sp.num_free_dofs for sp in spaces), sum(m.nnz for m in As)
    ils = fb.IterativeLinearSolver(); ils.set_type_from_string("CG")
    if multi:
        dinvs = [sp.create_dof_values() for sp in spaces]
        xs = [sp.create_dof_values() for sp in spaces]
    else:
        M = fb.JacobiPreconditioner(); M.create(op); M.symbolic_setup()
        ils.set_operators(A, M)
        x = space.create_dof_values()

    def sync_all():
        fb._lib.check(fb._lib.lib().fb2_device_synchronize())
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
            fb.jacobi_setup_multi(As, dinvs)
            t_tot, t_spmv = ils.iterate_multi(As, dinvs, bs, xs, a.cg_iters)
            return t_asm, t_tot, t_spmv
        op.force_compute()                                  # re-assembly into the fixed pattern every step
        t_asm = op.last_phase_ms("integrate")
        M.numerical_setup()
        t_tot, t_spmv = ils.iterate(op.get_translation(), x, a.cg_iters)
        return t_asm, t_tot, t_spmv

    for _ in range(a.warmup):
        step()
    sync_all()
    try:
        t_map_ms = op.last_phase_ms("scatter_map")
    except Exception:
        t_map_ms = None
    asm_ms, cg_ms, spmv_ms = [], [], []
    with ClockSampler(local) as clk:
        t0 = time.perf_counter()
        for _ in range(a.steps):
            ta, tt, ts = step()
            asm_ms.append(ta); cg_ms.append(tt); spmv_ms.append(ts)
        sync_all()
        wall = time.perf_counter() - t0
    # device-time per step (events on the library stream), max over ranks
    dev = torch.tensor([np.mean(asm_ms), np.mean(cg_ms), np.mean(spmv_ms) / a.cg_iters, wall / a.steps * 1e3],
                       dtype=torch.float64, device="cuda")
    tot = torch.tensor([float(n_loc), float(nnz_loc)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(dev, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    asm_t, cg_t, spmv_t, wall_t = [float(v) for v in dev.cpu()]
    n_glob_rows, nnz_glob = [float(v) for v in tot.cpu()]
    n_free_global = (a.order * cells - 1) ** 3
    bytes_all = spmv_bytes(n_glob_rows, nnz_glob)
    value = bytes_all / (spmv_t * 1e-3) / 1e9
    asm_mdofs = n_free_global / (asm_t * 1e-3) / 1e6

    # e2e: the same PCG iterations through the C-ABI call with HOST buffers (H2D of b, D2H of x inside the timed region)
    ils.set_parameters_from_pl({fb.ILS_RTOL: 0.0, fb.ILS_ATOL: -1.0, fb.ILS_MAX_NUM_ITERATIONS: a.cg_iters})
    ils.check_every = a.cg_iters
    ts = []
    if multi:
        b_hosts = [torch.from_numpy(o.get_translation().get_entries()).pin_memory() for o in ops]
        x_hosts = [torch.empty(sp.num_free_dofs, dtype=torch.float64).pin_memory() for sp in spaces]
        bs = [sp.create_dof_values() for sp in spaces]
        for k in range(3):
            sync_all()
            t0 = time.perf_counter()
            for bv, bh in zip(bs, b_hosts):
                bv.set_entries(bh.numpy())                  # H2D
            ils.apply_multi(As, dinvs, bs, xs)
            for xv, xh in zip(xs, x_hosts):
                fb._lib.check(fb._lib.lib().fb2_vec_get(xv.h, xh.data_ptr()))   # D2H
            ts.append(time.perf_counter() - t0)
    else:
        b_host = torch.empty(n_loc, dtype=torch.float64).pin_memory()
        b_host.copy_(torch.from_numpy(op.get_translation().get_entries()))
        x_host = torch.empty(n_loc, dtype=torch.float64).pin_memory()
        for k in range(3):
            sync_all()
            t0 = time.perf_counter()
            ils.apply(b_host.numpy(), x_host.numpy())
            fb._lib.check(fb._lib.lib().fb2_device_synchronize())
            ts.append(time.perf_counter() - t0)
    te = torch.tensor([min(ts[1:])], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e = {"value": a.cg_iters * bytes_all / float(te.item()) / 1e9, "unit": UNIT, "h2d_bytes_per_step": int(n_glob_rows * 8),
           "d2h_bytes_per_step": int(n_glob_rows * 8),
           "what": ("fb2_pcg_solve_host" if not multi else "fb2_vec_set + fb2_pcg_solve_multi + fb2_vec_get") +
                   " on every rank: H2D b + %d PCG iterations + D2H x (max over ranks)" % a.cg_iters}

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
    # DRAM bytes per SpMV launch from the committed ncu capture of this workload (profiles/r1_spmv_traffic.json), else null
    traffic = None
    try:
        tr = json.load(open(os.path.join(ROOT, "profiles", "r1_spmv_traffic.json")))
        key = "Q%d_%d_parts%d" % (a.order, cells, world)
        if key in tr:
            traffic = tr[key]["dram_bytes_per_launch"]
    except Exception:
        pass
    per_gpu_bytes = spmv_bytes(n_loc, nnz_loc)
    achieved = per_gpu_bytes / (np.mean(spmv_ms) / a.cg_iters * 1e-3) / 1e9
    asm_bytes = nnz_loc * 8 + n_loc * 8
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": asm_t + cg_t, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": "3D Poisson Q%d hex %d^3 cells (%d free DOFs, %d nnz), u=x+y+z Dirichlet, Jacobi-PCG; "
                                   "step = integrate+assemble + Jacobi setup + %d PCG iterations"
                                   % (a.order, cells, n_free_global, int(nnz_glob), a.cg_iters),
                       "parts": list(parts), "subdomains_per_gpu": ppr, "l2": "inputs larger than L2 (no flush needed)",
                       "scatter": a.scatter},
            "assembly_mdof_per_s": asm_mdofs, "assembly_ms": asm_t, "pcg_ms_per_iteration": cg_t / a.cg_iters,
            "spmv_ms": spmv_t, "setup_ms": {"fe_space": t_space * 1e3, "csr_pattern": t_pattern_ms, "scatter_map": t_map_ms},
            "wall_ms_per_step": wall_t,
            "roofline": {"bound": "hbm", "kernel": "k_spmv_stream<true,false> (TMA-streamed CSR row tiles, short-lived CTAs over 32 tiles, fused <Ap,p>)" if a.spmv_variant == 0 else "k_spmv<true> (warp-per-row)", "achieved": achieved,
                         "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                         "algorithmic_bytes_per_launch": per_gpu_bytes,
                         "frac_of_nominal_8000_GBps": achieved / 8000.0,
                         "peak_source": "measured (MEASURED_PEAKS.json hbm_gbs)" if peaks else "fallback 6650"},
            "roofline_assembly": {"bound": "hbm", "kernel": "k_assemble_fast<2,1,atomic>" if (a.asm_variant == 0 and a.order <= 2) else "k_assemble_generic", "achieved": asm_bytes / (np.mean(asm_ms) * 1e-3) / 1e9,
                                  "peak": peak, "unit": "GB/s", "frac": asm_bytes / (np.mean(asm_ms) * 1e-3) / 1e9 / peak},
            # BLAS-1 traffic of one PCG iteration: the reference's unfused vector TBPs move 18 n doubles, the fused
            # kernels here 11 n (x/r update with four partial dots: 4 reads + 2 writes + dinv; p update: 3 reads + 1 write)
            "pcg_vector_bytes_per_iteration": {"reference_unfused": int(18 * 8 * n_glob_rows), "fused": int(11 * 8 * n_glob_rows),
                                               "measured_ms": (cg_t / a.cg_iters) - spmv_t},
            "clocks": clk.summary(), "e2e": e2e,
            # kernels of libfempar_b200.so per step: Dirichlet interpolation + memsets/assembly (3), Jacobi set-up (2), PCG set-up
            # (9), then per PCG iteration: serial 6 (SpMV, 2 partial reductions, x/r update, scalars, p update); par 14 (two
            # SpMV launches, 4 pack/unpack kernels, interface dot, reductions, updates; NCCL's own kernels not counted)
            "gpu_launches": a.steps * ((3 + 2) * ppr + 9 + (6 if (world == 1 and not multi) else 14 * ppr) * a.cg_iters)}
    if not a.no_cpu_baseline and world == 1:       # reported on rank 0 at N=1 only
        r = cpu_reference_run(a.cpu_cells, a.order, a.cg_iters, 1, 0)
        line["cpu_baseline"] = {"value": r["spmv_gbs"], "unit": UNIT, "cores": r["cores"], "kind": "port", "sample": r["sample"],
                                "assembly_mdof_per_s": r["asm_mdofs"]}
    sys.stdout.flush()
    os.dup2(saved_stdout, 1)
    print(json.dumps(line), flush=True)
    os.dup2(2, 1)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
