import sys, json
sys.path.insert(0, '.')
import bench
bench.host_threads()
print(json.dumps(bench.b_ref_baselines()))
