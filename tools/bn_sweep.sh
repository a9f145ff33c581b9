# GPU tool: N-tile (BN) choices of the RAFT update block against bench throughput (three shards per GPU)
run() { env "$@" python bench.py --steps 30 --warmup 4 --no-train --no-cpu-baseline 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('%s value %.1f e2e %.1f single %.1f' % (sys.argv[1], d['value'], d['e2e']['value'], d['single_shard']['value']))" "$*"; }
run ZT_X=0
run ZT_BN_C2=96
run ZT_BN_ENC=128
run ZT_BN_Q=128
run ZT_BN_C2=96 ZT_BN_ENC=128
run ZT_SHARDS_PER_GPU=4
