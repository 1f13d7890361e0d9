import faulthandler, os, sys
faulthandler.dump_traceback_later(25, exit=True)
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["HATB200_CUDA_GRAPH"] = "0"
import torch
from super_resolution_b200 import HAT
cfg = dict(upscale=4, in_chans=3, img_size=64, window_size=16, compress_ratio=3, squeeze_factor=30, conv_scale=0.01,
           overlap_ratio=0.5, img_range=1., depths=[2, 2], embed_dim=48, num_heads=[6, 6], mlp_ratio=2,
           upsampler='pixelshuffle', resi_connection='1conv')
H = int(sys.argv[1]) if len(sys.argv) > 1 else 16
net = HAT(**cfg).eval().cuda()
net.engine().set_option("sync_debug", 1)
x = torch.rand(1, 3, H, H).bfloat16().cuda()
print("start", flush=True)
y = net(x)
torch.cuda.synchronize()
print("done", float(y.float().mean()), flush=True)
