#!/bin/bash
mkdir -p gpurun_out
timeout 200 python tools/host_profile.py --views 1 > gpurun_out/r02_host_profile_1view.log 2>&1; head -3 gpurun_out/r02_host_profile_1view.log
