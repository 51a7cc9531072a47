import json,sys
for r in json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])["results"]: print(r["shape"], r["residual"], round(r["fused_us"],1), r["aten_us"] and round(r["aten_us"],1))
