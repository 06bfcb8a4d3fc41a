"""How sensitive is one ResNet-56 training step (random init, batch 10) to arithmetic?  Gradient cosine between the plain
PyTorch fp64 / fp32 back-ends and the native tf32 / bf16 paths, for depth [1,1,1] .. [6,6,6]."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from fedml_b200.native.build import build
build()
from fedml_b200.models.batched import NativeBackend, TorchBackend, make_batched
from fedml_b200.models.cv import Bottleneck, ResNet
from fedml_b200.parallel.arena import ArenaLayout
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
DEV = "cuda:0"
from fedml_b200.models.cv import BasicBlock
for depth in ([1, 1, 1], [3, 3, 3], [6, 6, 6], "resnet18"):
    torch.manual_seed(0)
    C, B = 8, 10
    if depth == "resnet18":
        from fedml_b200.models import resnet18
        model = resnet18(10)
    else:
        model = ResNet(Bottleneck, depth, num_classes=10)
    layout = ArenaLayout(model)
    stack = torch.stack([layout.flatten(model.state_dict()) for _ in range(C)]).to(DEV)
    x = torch.randint(0, 256, (C, B, 32, 32, 3), dtype=torch.uint8, device=DEV)
    y = torch.randint(0, 10, (C, B), device=DEV)
    active = torch.ones(C, device=DEV)
    outs = {}
    for name, be in (("fp64", TorchBackend(torch.float64)), ("fp32", TorchBackend(torch.float32)),
                     ("lib_tf32", TorchBackend(torch.float32)),      # cuBLAS with TF32 allowed: what a library TF32 path does
                     ("tf32", NativeBackend(torch.float32)), ("bf16", NativeBackend(torch.bfloat16))):
        torch.backends.cuda.matmul.allow_tf32 = name == "lib_tf32"
        st, gst = stack.clone(), torch.zeros_like(stack)
        if name == "fp64":
            st, gst = st.double(), gst.double()
        W, G = layout.stacked_views(st), layout.stacked_views(gst)
        bm = make_batched(model, be)
        loss = bm.train_step(x, y, W, G, active.to(st.dtype) if name == "fp64" else active)
        outs[name] = (loss.double().clone(), gst[:, :layout.P_train].double().clone())
    ref = outs["fp64"][1]
    for name in ("fp32", "lib_tf32", "tf32", "bf16"):
        g = outs[name][1]
        cos = torch.nn.functional.cosine_similarity(g, ref, dim=1)
        rel = (g - ref).norm(dim=1) / ref.norm(dim=1)
        print(f"depth {depth} {name}: cos min {float(cos.min()):.6f} rel max {float(rel.max()):.5f} loss diff {float((outs[name][0]-outs['fp64'][0]).abs().max()):.2e}")
