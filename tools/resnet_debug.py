import sys, torch
sys.path.insert(0, "/root/repo")
import dtg_b200  # noqa: F401
from dtg_b200.models.resnet import tiny_resnet
from dtg_b200.models.resnet_native import NativeResNet
def rel_err(out, ref):
    out, ref = out.float(), ref.float()
    return ((out - ref).abs().max() / (ref.abs().max() + 1e-6)).item()
torch.manual_seed(0)
ref = tiny_resnet()
B = int(sys.argv[1]) if len(sys.argv) > 1 else 16
nat = NativeResNet(ref.cfg, batch_size=B, device="cuda")
params = {k: v.cuda() for k, v in ref.init_params(seed=3).items()}
pr = {k: (v.bfloat16().float() if v.dim() == 2 else v) for k, v in params.items()}
x, y = nat.synthetic_batch()
loss_ref, g_ref = ref.forward_backward(pr, (x.bfloat16().float(), y))
loss, g = nat.forward_backward(params, (x, y))
print("loss", loss.item(), loss_ref.item())
for k in g_ref:
    cos = torch.nn.functional.cosine_similarity(g[k].float().flatten(), g_ref[k].float().flatten(), dim=0).item()
    print("cos %.4f " % cos, end="")
    print("%-60s err %.3f  |ref| %.3e |nat| %.3e" % (k, rel_err(g[k], g_ref[k]), g_ref[k].abs().max().item(), g[k].float().abs().max().item()))
