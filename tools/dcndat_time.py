import sys, os, json, torch
sys.path.insert(0, os.getcwd()); sys.path.insert(0, 'tests')
import bench
class A: pass
a = A(); a.steps = 10; a.tf32 = 1
torch.backends.cudnn.allow_tf32 = True; torch.backends.cudnn.benchmark = True
dev = torch.device('cuda', 0)
print(json.dumps(bench.dcndat_1080p(a, dev, 1080, 1920, 1, None, with_library_baseline=False)))
