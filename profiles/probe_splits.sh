# developer probe: FlowLM chain time against the split-K factor of each decode GEMM
run() { echo -n "SPLITS=$1: "; PTTS_SPLITS="$1" PTTS_HEAD=0 PTTS_PROFILE_PART=3 python profiles/probe_concurrent_chains.py 1 2>&1 | tail -1 | cut -d, -f2; }
run ""
run "flow.ffn1:2"
run "flow.ffn1:1"
run "flow.qkv:2"
run "flow.ffn2:4"
run "flow.out_proj:4"
run "flow.out_proj:2,flow.ffn2:2"
run "head.mlp0:4,head.mlp2:4"
run "head.mlp0:2,head.mlp2:2"
run "head.adaln_mod:1,head.cond_embed:4"
