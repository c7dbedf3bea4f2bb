import sys, json
for l in sys.stdin:
    try:
        d = json.loads(l)
    except Exception:
        if "rror" in l:
            print(l.rstrip()[:200])
        continue
    print(d.get("Hkv"), d.get("S"), d.get("causal"), round(d.get("tflops", 0), 1))
