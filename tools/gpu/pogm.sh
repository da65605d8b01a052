python tests/devtools/profile_cases.py ogm > gpurun_out/pogm_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -k regex:ogm_ --csv --log-file gpurun_out/pogm_launches.csv python tests/devtools/profile_cases.py ogm > gpurun_out/pogm_ncu.log 2>&1
tail -3 gpurun_out/pogm_plain.log
