python - <<'PY'
import cProfile, pstats, torch, time
import bench
for _ in range(3): bench.api_call("ref", 1, None)
torch.cuda.synchronize()
t0=time.perf_counter()
for _ in range(10): bench.api_call("ref", 1, None)
print("wall per call ms", (time.perf_counter()-t0)/10*1e3)
pr = cProfile.Profile(); pr.enable()
for _ in range(20): bench.api_call("ref", 1, None)
pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(14)
PY
