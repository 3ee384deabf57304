#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -c "
import cProfile, pstats, sys, io
sys.argv = ['bench.py', '--workload', 'cfg5']
import runpy
pr = cProfile.Profile()
pr.enable()
try:
    runpy.run_path('bench.py', run_name='__main__')
except SystemExit:
    pass
pr.disable()
s = io.StringIO()
pstats.Stats(pr, stream=s).sort_stats('cumulative').print_stats(45)
open('gpurun_out/r02_cfg5_cprofile.txt', 'w').write(s.getvalue())
" > gpurun_out/r02_cfg5_prof.json 2>/dev/null
grep -A60 "cumulative" gpurun_out/r02_cfg5_cprofile.txt | cut -c1-150 | head -70
