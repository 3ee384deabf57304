#!/bin/bash
# What each group of Mimi-branch kernels costs the FULL frame step (FlowLM || Mimi): drop the group's launches
# (PTTS_SKIP_TAGS, garbage results) and time the 64-stream loop.  Compare with the group's own machine time
# (profiles/r02_mimi_kernel_times.txt) - a group that costs the step more than it runs is hurting the FlowLM chain.
run() { echo -n "skip=[$1]: "; PTTS_SKIP_TAGS="$1" python profiles/probe_concurrent_chains.py 1 2>&1 | tail -1; }
run ""
run "seanet.stage2_fused"
run "seanet.resblock_fused"
run "mimi.attn"
run "mimi.qkv,mimi.out_proj,mimi.ffn1,mimi.ffn2"
run "seanet.conv0,seanet.convtr0,seanet.convtr1"
run "mimi.ln,mimi.embed_upsample,seanet.hist_gather,seanet.hist_scatter,seanet.stage"
run "seanet.stage2_fused,seanet.resblock_fused,seanet.conv0,seanet.convtr0,seanet.convtr1"
run "mimi.attn,mimi.qkv,mimi.out_proj,mimi.ffn1,mimi.ffn2,mimi.ln,mimi.embed_upsample"
run "seanet.stage2_fused,seanet.resblock_fused,seanet.conv0,seanet.convtr0,seanet.convtr1,mimi.attn,mimi.qkv,mimi.out_proj,mimi.ffn1,mimi.ffn2,mimi.ln,mimi.embed_upsample,seanet.hist_gather,seanet.hist_scatter,seanet.stage"
run ""
