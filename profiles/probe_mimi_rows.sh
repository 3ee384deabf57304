# developer probe: Mimi-only frame time against the k-block threshold of the persistent row-major GEMM and its smem budget
for kb in 8 4 3 2 1; do echo -n "ROWS_MINKB=$kb: "; PTTS_TC_ROWS_MINKB=$kb PTTS_PROFILE_PART=4 python profiles/probe_concurrent_chains.py 1 2>&1 | tail -1 | cut -d, -f2; done
for kb in 8 1; do echo -n "ROWS_MINKB=$kb SMEM=200: "; PTTS_TC_ROWS_SMEM_KB=200 PTTS_TC_ROWS_MINKB=$kb PTTS_PROFILE_PART=4 python profiles/probe_concurrent_chains.py 1 2>&1 | tail -1 | cut -d, -f2; done
for g in 8; do echo -n "MIMI_GROUP=$g: "; PTTS_MIMI_GROUP=$g PTTS_PROFILE_PART=4 python profiles/probe_concurrent_chains.py 1 2>&1 | tail -1 | cut -d, -f2; done
