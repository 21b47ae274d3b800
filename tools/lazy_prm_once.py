import sys; sys.path.insert(0,'/root/repo')
import bench, json
print(json.dumps(bench.measure_lazy_prm_c2(), indent=1))
