run() { echo -n "SPLITS=$1: "; PTTS_SPLITS="$1" python bench.py --workload cfg3 --steps 32 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], {k: v for k, v in d['kernel_ms_share'].items() if k.startswith('flow.')})"; }
run ""
run "flow.ffn1:2,flow.ffn2:2,flow.qkv:2,flow.out_proj:2"
run "flow.ffn1:1,flow.ffn2:2,flow.qkv:1,flow.out_proj:2"
run "flow.ffn1:1,flow.ffn2:4,flow.qkv:1,flow.out_proj:4"
run "flow.ffn1:2,flow.ffn2:4,flow.qkv:2,flow.out_proj:8"
