#!/bin/bash
# par apply: sequential vs overlapped halo exchange, and the chunk size of the short-lived interior CTAs (run under torchrun)
N=${N:-2}
run() { env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 2 --warmup 2 --no-cpu-baseline 2>/dev/null | grep '^{' | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('value %.0f GB/s spmv+halo %.3f ms pcg/it %.3f ms asm %.2f ms'%(d['value'],d['spmv_ms'],d['pcg_ms_per_iteration'],d['assembly_ms']))"; }
echo "N=$N sequential:"; run FB2_PAR_OVERLAP=0
for c in 16 64 256; do echo "N=$N overlap chunk=$c:"; run FB2_PAR_OVERLAP=1 FB2_PAR_CHUNK=$c; done
