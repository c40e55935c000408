import sys
sys.path[:0]=['/root/repo/pytorch-distributed-template_b200','/root/repo']
import torch, bench
d=torch.device('cuda')
for m in (0,1,2,3,4): print('mode',m, bench.measure_fma_peak(d,m))
