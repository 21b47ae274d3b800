import sys; sys.path.insert(0,'/root/repo')
import bench, json
print(json.dumps(bench.measure_smoothing_c5(), indent=1))
print(json.dumps(bench.measure_rrt_batch(1, 0, None), indent=1))
print(json.dumps(bench.measure_rrt_batch(1, 0, None, total_queries=65536), indent=1))
