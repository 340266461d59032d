#!/bin/bash
# on the GPU box: bench every scratch/variants/libvslam_*.so (kernel ms per step + value)
cd /root/repo
cp visual-slam_b200/orbslam2/libvslam_b200.so /tmp/orig.so
for so in scratch/variants/libvslam_*.so; do
  cp $so visual-slam_b200/orbslam2/libvslam_b200.so
  python bench.py --steps 10 --warmup 3 --no-cpu 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); k=d['kernel_ms_per_step']
print('$so'.split('_')[-1], 'value %.0f e2e %.0f levels %.3f' % (d['value'], d['e2e']['value'], sum(k['level%d'%i] for i in range(8))), ' '.join('%.3f'%k['level%d'%i] for i in range(8)), 'desc %.3f knn %.3f' % (k['describe'], k['knn2']))"
done
cp /tmp/orig.so visual-slam_b200/orbslam2/libvslam_b200.so
