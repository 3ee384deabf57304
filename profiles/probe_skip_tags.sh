run() { echo -n "skip=$1 head=$2 part=$3: "; PTTS_SKIP_TAGS="$1" PTTS_HEAD=$2 PTTS_PROFILE_PART=$3 python profiles/probe_concurrent_chains.py 1 2>&1 | tail -1 | cut -d, -f2; }
run "" 2 3
run "flow.attn_decode" 2 3
run "flow.qkv" 2 3
run "flow.out_proj" 2 3
run "flow.ffn1" 2 3
run "flow.ffn2" 2 3
run "flow.qkv,flow.out_proj,flow.ffn1,flow.ffn2" 2 3
run "head.cluster" 2 3
run "head.cond_embed,head.adaln_mod" 2 3
run "flow.embed,flow.final_norm_eos,flow.ln" 2 3
run "" 0 3
run "head.mlp0,head.mlp2" 0 3
run "head.ln_modulate" 0 3
run "flow.attn_decode,flow.qkv,flow.out_proj,flow.ffn1,flow.ffn2,head.cluster,head.cond_embed,head.adaln_mod" 2 3
