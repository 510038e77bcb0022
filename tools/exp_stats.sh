#!/bin/bash
CUDASW_NVCC_EXTRA="-DCUSW_STEP_STATS" python gridss_b200/build.py --force > /dev/null 2>&1
python tools/_stats_run.py 2>&1 | grep -E "cfg|step stats"
python gridss_b200/build.py --force > /dev/null 2>&1
