import json,sys
l=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1], round(l["value"],1), round(l["ms_per_step"],3), "e2e", round(l["e2e"]["value"],1),
      "store", round(l.get("f1_store_path", {}).get("value", 0), 1),
      "train it/s", round(l.get("train_step", {}).get("it_per_s", 0), 1), "ms", round(l.get("train_step", {}).get("ms_per_step", 0), 3))
print("  "+" ".join(f"{k}={v['ms']}" for k,v in l["roofline"]["stages"].items()))
