for f in 1 0 1 0; do
echo "== GBL_XCHG_FUSED=$f"
GBL_XCHG_FUSED=$f timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 20 --warmup 3 --no-secondary 2>/dev/null | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('value %.1f ms %.3f collective %.3f e2e %.1f'%(d['value'],d['ms_per_step'],d['roofline']['stage_share']['collective'],d['e2e']['value']))
"
done
