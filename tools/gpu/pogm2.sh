python tests/devtools/profile_cases.py ogm > gpurun_out/pogm_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:ogm_apply -s 1 -c 1 -o gpurun_out/pogm_apply python tests/devtools/profile_cases.py ogm > gpurun_out/pogm_ncu.log 2>&1
tail -2 gpurun_out/pogm_ncu.log
