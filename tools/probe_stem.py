"""GPU probe: cuDNN time of the ResNet stem convolution (7x7/2, 3->64, N=256, 224x224, bf16 NHWC) in
several equivalent formulations, forward and weight-gradient."""
import torch
import torch.nn.functional as F

torch.backends.cudnn.benchmark = True
dev = "cuda"


def timeit(fn, reps=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def case(name, cin, k, stride, pad, hw, cout=64):
    x = torch.randn(256, cin, hw, hw, device=dev, dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last)
    w = torch.randn(cout, cin, k, k, device=dev, dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last)
    y = F.conv2d(x, w, None, stride, pad)
    gy = torch.randn_like(y)
    f = timeit(lambda: torch.ops.aten.convolution(x, w, None, [stride] * 2, [pad] * 2, [1, 1], False, [0, 0], 1))
    b = timeit(lambda: torch.ops.aten.convolution_backward(gy, x, w, None, [stride] * 2, [pad] * 2, [1, 1], False,
                                                           [0, 0], 1, [False, True, False]))
    print(f"{name:34s} out {tuple(y.shape)}  fwd {f:.3f} ms  wgrad {b:.3f} ms", flush=True)


case("7x7/2 C=3 (reference layout)", 3, 7, 2, 3, 224)
case("7x7/2 C=4 (zero-padded channel)", 4, 7, 2, 3, 224)
case("7x7/2 C=8", 8, 7, 2, 3, 224)
case("space-to-depth 4x4/1 C=12", 12, 4, 1, 0, 115)   # 224+6 -> 230 -> /2 = 115 ; 115-4+1 = 112
case("space-to-depth 4x4/1 C=16", 16, 4, 1, 0, 115)
# other layers for context
case("3x3/1 C=64 56x56", 64, 3, 1, 1, 56)
case("1x1/1 C=256->64 56x56", 256, 1, 1, 0, 56)
x = torch.randn(256, 64, 112, 112, device=dev, dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last)
print("maxpool fwd ms", timeit(lambda: torch.ops.aten.max_pool2d_with_indices(x, [3, 3], [2, 2], [1, 1], [1, 1], False)))
