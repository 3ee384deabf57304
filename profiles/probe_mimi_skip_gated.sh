#!/bin/bash
# As probe_mimi_skip.sh, with the head-gated loop: what the remaining (ungated) Mimi kernel groups cost the full step
export PTTS_HEAD_GATE=1
run() { echo -n "skip=[$1]: "; PTTS_SKIP_TAGS="$1" python profiles/probe_concurrent_chains.py 1 2>&1 | tail -1; }
run ""
run "mimi.ln"
run "mimi.attn"
run "mimi.qkv,mimi.out_proj"
run "mimi.ffn1,mimi.ffn2"
run "mimi.embed_upsample,seanet.stage,seanet.hist_gather,seanet.hist_scatter"
run "seanet.conv0"
run "seanet.convtr0,seanet.convtr1"
run "seanet.resblock_fused"
run "seanet.stage2_fused"
run "mimi.ln,mimi.attn,mimi.qkv,mimi.out_proj,mimi.ffn1,mimi.ffn2,mimi.embed_upsample"
run "seanet.stage2_fused,seanet.resblock_fused,seanet.conv0,seanet.convtr0,seanet.convtr1,mimi.attn,mimi.qkv,mimi.out_proj,mimi.ffn1,mimi.ffn2,mimi.ln,mimi.embed_upsample,seanet.hist_gather,seanet.hist_scatter,seanet.stage"
run ""
