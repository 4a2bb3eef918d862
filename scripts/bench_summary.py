import json,sys
d=json.load(open(sys.argv[1]))
print("ms/step",round(d["ms_per_step"],1),"iters",d["lm_iters_per_step"],"pcg",d["pcg_iters_per_step"],"cost",d["final_cost"],"value",d["value"]/1e6,"e2e",d["e2e"]["value"]/1e6, d["status"])
for k,v in d["roofline"]["kernels"].items(): print(f'{k:24s} {v["launches"]:6d} {v["ms"]:9.3f} ms {1e3*v["ms"]/v["launches"]:8.1f} us {v["GBps"]:7.0f} GB/s')
