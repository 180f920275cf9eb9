import json,sys
d=json.load(open(sys.argv[1]))
print(sys.argv[1], "value %.0f ms/step %.3f"%(d["value"], d["ms_per_step"]), d["extra"])
print("   "+"  ".join("%s %.2f"%(k.replace("fwd_","f").replace("inv_","i"),v["ms"]) for k,v in d["kernels"].items() if v["ms"]>0.2))
