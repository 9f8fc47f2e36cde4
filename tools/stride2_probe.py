"""How fast can the library do the full-resolution stride-2 conv of the encoder (72 -> 72, 1088x1920 -> 544x960)?"""
import torch, torch.nn.functional as F
dev = torch.device("cuda:0")
torch.backends.cudnn.allow_tf32 = True
x = torch.randn(1, 72, 1088, 1920, device=dev)
w = torch.randn(72, 72, 3, 3, device=dev) * 0.05
b = torch.randn(72, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
def timeit(fn, n=10):
    for _ in range(3): fn()
    ts = []
    for _ in range(n):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return sorted(ts)[len(ts) // 2]
for bench in (False, True):
    torch.backends.cudnn.benchmark = bench
    for name, xx, ww in (("nchw", x, w), ("nhwc", x.contiguous(memory_format=torch.channels_last), w.contiguous(memory_format=torch.channels_last))):
        print(f"benchmark={bench} {name}: {timeit(lambda: F.conv2d(xx, ww, b, stride=2, padding=1)):.3f} ms", flush=True)
xc, wc = x.contiguous(memory_format=torch.channels_last), w.contiguous(memory_format=torch.channels_last)
print(f"nhwc x, nchw w, no bias: {timeit(lambda: F.conv2d(xc, w, None, stride=2, padding=1)):.3f} ms")
print(f"nhwc x, nhwc w, no bias: {timeit(lambda: F.conv2d(xc, wc, None, stride=2, padding=1)):.3f} ms")
print(f"nhwc x, nchw w, bias: {timeit(lambda: F.conv2d(xc, w, b, stride=2, padding=1)):.3f} ms")
for (c, hh, ww_) in ((72, 544, 960), (72, 272, 480), (72, 136, 240)):
    xl = torch.randn(1, c, hh, ww_, device=dev).contiguous(memory_format=torch.channels_last)
    print(f"{hh}x{ww_}: nchw w {timeit(lambda: F.conv2d(xl, w, None, stride=2, padding=1)):.4f} ms, nhwc w {timeit(lambda: F.conv2d(xl, wc, None, stride=2, padding=1)):.4f} ms")
# space-to-depth form: a 2x2 stride-1 conv over 288 channels at half resolution (zero weights where the 3x3 has no tap)
xs = F.pixel_unshuffle(F.pad(x, (1, 1, 1, 1)), 2).contiguous(memory_format=torch.channels_last)   # (1, 288, 545, 961)
w2 = torch.zeros(72, 72, 4, 4, device=dev); w2[:, :, :3, :3] = w
w2 = w2.view(72, 72, 2, 2, 2, 2).permute(0, 1, 3, 5, 2, 4).reshape(72, 288, 2, 2).contiguous(memory_format=torch.channels_last)
ref = F.conv2d(x, w, b, stride=2, padding=1)
y = F.conv2d(xs, w2, b)
print("s2d shape", tuple(y.shape), "max diff", (y - ref).abs().max().item())
print(f"s2d conv alone: {timeit(lambda: F.conv2d(xs, w2, b)):.3f} ms")
