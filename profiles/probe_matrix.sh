# developer probe: frame time (ms per 64-stream frame, generate_frames) over the kernel-form switches
run() { echo -n "HEAD=$1 ATTN=$2 LNFOLD=$3 PART=$4: "; PTTS_HEAD=$1 PTTS_ATTN=$2 PTTS_LNFOLD=$3 PTTS_PROFILE_PART=$4 python profiles/probe_concurrent_chains.py 1 2>&1 | tail -1 | cut -d, -f2; }
for part in 3 0; do
run 0 1 1 $part
run 2 1 1 $part
run 0 2 1 $part
run 0 1 0 $part
run 0 0 0 $part
done
