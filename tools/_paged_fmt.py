import sys,json
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: print(l.rstrip()); continue
    print(d["shape"], d["flat"], round(d["paged_us"],1), round(d["frac_hbm"],3), d.get("host_lengths_would_pick"), d.get("contiguous_us"), d.get("flash_attn2_us"))
