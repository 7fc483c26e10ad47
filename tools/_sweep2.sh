# developer sweep: occupancy / tail / budget sensitivity of the tiled kNN kernel on the KITTI scan
for c in 7 6 5 4 3; do echo -n "CTAS=$c  "; PCP_TILE_CTAS=$c python tools/dev_sweep.py kitti | cut -c1-160; done
for t in 0.02 0.1 0.2 0.35; do echo -n "TAIL=$t  "; PCP_TILE_TAIL=$t python tools/dev_sweep.py kitti | cut -c1-160; done
for s in 16 24 48; do echo -n "SCANS=$s "; PCP_TILE_MAX_SCANS=$s python tools/dev_sweep.py kitti | cut -c1-160; done
for c in 5 4 3; do echo -n "uni CTAS=$c  "; PCP_TILE_CTAS=$c python tools/dev_sweep.py uni | cut -c1-160; done
