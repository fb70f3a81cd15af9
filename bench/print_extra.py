import json, sys
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print(sys.argv[1] if len(sys.argv) > 1 else "", "terasort", round(d["value"], 1), "join_ms", round(d["extra"]["broadcast_join"]["ms_per_step"], 4),
      "kmeans_ms", round(d["extra"]["kmeans"]["value"] * 1e3, 4), "groupby_Grows", round(d["groupby_agg"]["value"] / 1e9, 1))
