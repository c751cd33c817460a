for r in 2000 3500 6000 8000 12000 20000; do echo "range $r"; PDG_M3_RANGE=$r python tools/profile_step.py 100 | grep -o "'ms_hist.*'ms_pass': \[[^]]*\]"; done
