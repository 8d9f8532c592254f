for a in "" "--causal" "--fp8" "--legacy"; do FA100_LIB=libfa100_prof.so timeout 120 python tools/prof2.py $a; done 2>&1 | grep -v Warning
AB_ROUNDS=2 timeout 600 python tools/ab_time.py libfa100.so libfa100_p63.so libfa100_r192.so 2>&1
