"""Where the complete step's extra time goes: VAE encode (B = 8, 512x512), CLIP-L forward + backward (24 sequences)."""
import os, sys, time
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from stable_diffusion_training_b200 import clip as CL, lion_quant as LQ, training_utils as TU, vae as PV, _lib

dev = torch.device("cuda:0")
def timed(fn, n=5):
    for _ in range(2): fn()
    torch.cuda.synchronize(); e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n

vtree = LQ.unflatten_dict(PV.init_params_numpy(PV.SD_VAE, 2))
enc = PV.AutoencoderKL(PV.SD_VAE).encoder(vtree, dev)
px = torch.rand(8, 3, 512, 512, device=dev) * 2 - 1
n0 = _lib.lib().sdb_launch_count(); enc.encode(px); print("vae launches", _lib.lib().sdb_launch_count() - n0)
print("vae encode ms", timed(lambda: enc.encode(px)))
# per-stage
import types
x = torch.randn(8, 512, 512, 8, device=dev).to(torch.bfloat16)
from stable_diffusion_training_b200 import ops
x128 = torch.randn(8, 512, 512, 128, device=dev).to(torch.bfloat16)
print("  conv_in (tc, 8-channel pixels) ms", timed(lambda: ops.conv3x3(x, enc._conv_in_w, bias=enc.P["encoder/conv_in/bias"])))
print("  resnet 128@512 ms", timed(lambda: enc._resnet(x128, "encoder/down_blocks_0/resnets_0")))
print("  gn 128@512 ms", timed(lambda: enc._gn(x128, 8, "encoder/down_blocks_0/resnets_0/norm1", True)))
print("  conv 128->128@512 ms", timed(lambda: enc._conv(x128, "encoder/down_blocks_0/resnets_0/conv1")))
print("  down 128@512 ms", timed(lambda: enc._downsample(x128, "encoder/down_blocks_0/downsamplers_0/conv")))
x512 = torch.randn(8, 64, 64, 512, device=dev).to(torch.bfloat16)
print("  attention 512@64 ms", timed(lambda: enc._attention(x512, "encoder/mid_block/attentions_0")))
print("  conv_out simt ms", timed(lambda: enc._conv(x512, "encoder/conv_out")))
tparams = LQ.unflatten_dict({k: torch.from_numpy(v) for k, v in CL.init_params_numpy(CL.SD15_CLIP, 1).items()})
tstate = TU.TrainState.create(apply_fn=CL.CLIPTextModel(CL.SD15_CLIP).__call__, params=tparams, tx=TU.create_lion_optimizer(tparams), device=dev)
ids = torch.randint(0, 49408, (24, 77), device=dev, dtype=torch.int32)
dO = torch.randn(24 * 77, 768, device=dev).to(torch.bfloat16)
def clip_fb():
    out = tstate.net.forward(ids); out.g = dO; tstate.net.backward()
print("clip fwd+bwd ms (eager)", timed(clip_fb))
print("clip optimizer ms", timed(lambda: tstate.apply_gradients(grads=None)))
