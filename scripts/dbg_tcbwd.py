import sys, json
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import torch
import test_gpu_tc as T
out = {}
for name in ("cfg1", "cfg2"):
    for prec in ("fp32", "bf16x3", "bf16"):
        try:
            errs, _ = T._grad_errors(name, prec)
            top = sorted(errs.items(), key=lambda kv: -kv[1])[:8]
            out[f"{name}_{prec}"] = errs
            print(name, prec, top, flush=True)
            continue
        except Exception as e:
            out[f"{name}_{prec}"] = repr(e)
        json.dump(out, open("/root/repo/gpurun_out/dbg_tcbwd.json", "w"), indent=1)
        print(name, prec, out[f"{name}_{prec}"], flush=True)
