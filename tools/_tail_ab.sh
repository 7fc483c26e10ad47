#!/bin/sh
# developer sweep: KITTI step against the tail hand-over share, overflow kernel overlapped or serial
for ov in 1; do for t in 0.0 0.01 0.02 0.03 0.04 0.05 0.07; do
  echo "== overlap=$ov tail=$t"
  PCP_KNN_OVERLAP=$ov PCP_TILE_TAIL=$t python bench.py --steps 100 --warmup 10 --no-cpu-baseline --icp-pairs 0 --no-radius 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read());print(round(d['value']/1e6,1),'Mq/s step',round(d['ms_per_step'],4),'tile',round(d['roofline']['kernel_ms'],4),'frac',round(d['roofline']['frac'],4),'search',round(d['breakdown_ms']['knn'],4), 'lat', round(d['latency']['single_shot_ms'],4))"
done; done
