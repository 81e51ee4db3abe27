import os, sys, time
sys.path.insert(0, os.getcwd())
import numpy as np, torch, torch.distributed as dist
import fempar_b200 as fb
rank, world, local = int(os.environ.get("RANK",0)), int(os.environ.get("WORLD_SIZE",1)), int(os.environ.get("LOCAL_RANK",0))
torch.cuda.set_device(local); fb.init(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local)); fb.comm_init_from_torch(dist)
cells = int(sys.argv[1]); parts = tuple(int(v) for v in sys.argv[2].split(","))
part_ids = [rank + 1] if world > 1 else list(range(1, int(np.prod(parts)) + 1))
for pid in part_ids:
    s = fb.FeSpace([cells]*3, order=2, num_parts_x_dim=parts, part_id=pid)
    op = fb.FeAffineOperator(); op.create(fb.csr_format, [False], [True], [1], s, fb.DiscreteIntegration(fb.poisson, fb.func_linear))
    for k in range(3):
        op.compute()
    ph = lambda n: op.last_phase_ms(n)
    x, y = s.create_dof_values(), s.create_dof_values(); x.init(1.0)
    sp, ha = [], []
    for k in range(6):
        op.get_matrix().apply(x, y); sp.append(ph("spmv")); ha.append(ph("halo"))
    print("rank %d part %d n=%d nI=%d nG=%d nnz=%d pattern %.1f map %.1f ms (class %d, tables %d) integrate %.2f ms spmv %.3f halo %.3f R %d tiles %d cta/sm %d GB/s %.0f" % (
        rank, pid, s.num_free_dofs, s.num_interior_dofs, s.num_interface_dofs, op.get_matrix().nnz, ph("pattern"), ph("scatter_map"),
        ph("scatter_map_class_tables"), ph("scatter_map_tables"), ph("integrate"), min(sp), min(ha), ph("spmv_R"), ph("spmv_tiles"), ph("spmv_ctas_per_sm"),
        (op.get_matrix().nnz * 12 + s.num_free_dofs * 24) / min(sp) / 1e6), flush=True)
    s.free()
if world > 1:
    dist.destroy_process_group()
