import copy, sys, torch, torchvision
sys.path.insert(0, ".")
from garfield_b200.models.fused_resnet import fuse_resnet_blocks
torch.backends.cudnn.allow_tf32 = False; torch.backends.cudnn.deterministic = True
torch.manual_seed(0)
def grads(model, x, t):
    for p in model.parameters(): p.grad = None
    torch.nn.functional.cross_entropy(model(x), t).backward()
    return [p.grad.clone() for p in model.parameters()]
def report(tag, ga, gb, names):
    fa, fb = torch.cat([g.reshape(-1) for g in ga]), torch.cat([g.reshape(-1) for g in gb])
    cos = torch.nn.functional.cosine_similarity(fa, fb, dim=0).item()
    worst, wname = 0, None
    for n, a, b in zip(names, ga, gb):
        r = (a - b).abs().max().item() / (b.abs().max().item() + 1e-12)
        if r > worst: worst, wname = r, n
    print("%-28s cos %.6f  rel-L2 %.3e  worst per-param max-rel %.3e (%s)" % (tag, cos, ((fa - fb).norm() / fb.norm()).item(), worst, wname), flush=True)
ref = torchvision.models.resnet50(num_classes=10).cuda().to(memory_format=torch.channels_last).train()
names = [n for n, _ in ref.named_parameters()]
ours = copy.deepcopy(ref); fuse_resnet_blocks(ours)
x = torch.randn(25, 3, 32, 32, device="cuda").contiguous(memory_format=torch.channels_last)
t = torch.randint(0, 10, (25,), device="cuda")
g_ref = grads(ref, x, t)
ref2 = copy.deepcopy(ref)
g_pert = grads(ref2, x * (1 + 1e-6), t)       # sensitivity of the REFERENCE to a 1e-6 input perturbation
report("reference vs perturbed ref", g_pert, g_ref, names)
report("fused vs reference", grads(ours, x, t), g_ref, names)
# a single bottleneck block (well conditioned)
blk = copy.deepcopy(ref.layer2[1]); blk_f = copy.deepcopy(blk)
import types
from garfield_b200.models.fused_resnet import _bottleneck_forward
blk_f.forward = types.MethodType(_bottleneck_forward, blk_f)
xb = torch.randn(25, 512, 4, 4, device="cuda").contiguous(memory_format=torch.channels_last)
def bg(b):
    for p in b.parameters(): p.grad = None
    xi = xb.clone().requires_grad_(True)
    b(xi).square().mean().backward()
    return [p.grad.clone() for p in b.parameters()] + [xi.grad.clone()]
report("one block: fused vs ref", bg(blk_f), bg(blk), [n for n, _ in blk.named_parameters()] + ["input"])
