"""Random-gather microbenchmark + count kernel time on the human-scale filter for one L2 fetch granularity (env)."""
import json, os, subprocess, sys
out = subprocess.run([sys.executable, "bench.py", "--workload", sys.argv[1] if len(sys.argv) > 1 else "human", "--steps", "6", "--warmup", "3", "--no-cpu-baseline"],
                     capture_output=True, text=True).stdout
d = json.loads(out)
g = d["roofline"]["gather_microbench"]
print("L2_FETCH", os.environ.get("CHARON_L2_FETCH_BYTES", "default32"), "value %.3f M/s" % (d["value"] / 1e6), "count_ms %.3f" % d["roofline"]["avg_launch_ms"],
      "gather query %.0f GB/s sector %.0f GB/s frac %.3f" % (g["query_addressing_GBps"], g["pure_sector_GBps"], g["frac_of_achievable_gather"]))
