import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1], "value",round(d["value"],1),"step_ms",round(d["ms_per_step"],4),"e2e_ms",round(d["e2e"]["ms_per_step"],4),"k1_ms",round(d["roofline"]["avg_launch_ms"],4),"frac",round(d["roofline"]["frac"],3),"items",d["pool"]["n_work_items"])
