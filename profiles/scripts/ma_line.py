import sys, json
for l in sys.stdin:
    d = json.loads(l)
    print(sys.argv[1], d["envs"], d["l2_flushed_between_launches"], round(d["ms_step_kernel_median"]*1e3,1), "us", round(d["frac_algorithmic"],3))
