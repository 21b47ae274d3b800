import sys; sys.path.insert(0,'/root/repo')
import bench, json
for it in (20000, 200000):
    print(json.dumps(bench.measure_rrt_star(it)))
