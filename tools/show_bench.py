import json,sys
d=json.load(open(sys.argv[1]))
print(sys.argv[1], round(d["value"],1), round(d["ms_per_step"],3), round(d["e2e"]["value"],1))
print(' '.join(f'{k.replace("op_","").replace("linear_tc:","")}={v["ms_per_step"]:.3f}' for k,v in d["kernels"].items()))
