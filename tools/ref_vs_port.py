"""Build-container only (needs /root/reference, like tests/golden/make_golden.py; nothing on the GPU box runs this): the unmodified reference (torch_scatter stand-in) vs the oracle port,
config-4 sample (2 graphs x 100 particles, H = 256), train step on the host cores."""
import sys, time, torch
import os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import make_golden as mg
mg._install_shims()
from nn.state2state.aether import Aether
from oracle import aether_oracle as O
torch.set_num_threads(8)
D, N, B, H = 2, 100, 2, 256
torch.manual_seed(0)
model = Aether(input_size=2 * D, hidden_size=H, dropout_prob=0.0, num_dims=D, device="cpu")
p = {k: v.detach().clone().requires_grad_(True) for k, v in model.state_dict().items()}
x, v, q = O.synth_state(B, N, D, seed=1)
send, recv = O.fc_edges(N, B)
h, ea = O.caller_prep(x, v, q, send, recv)
tgt = x + 0.01
def ref_step():
    model.zero_grad()
    out = model(h, x, [send, recv], v, ea, q)
    torch.nn.functional.mse_loss(out, tgt).backward()
    return out
def port_step():
    for t in p.values(): t.grad = None
    out = O.aether_step(p, x, (send, recv), v, ea, q)
    torch.nn.functional.mse_loss(out, tgt).backward()
    return out
a, b = ref_step(), port_step()
print("max |ref - port| =", float((a - b).detach().abs().max()))
for name, fn in (("reference", ref_step), ("port", port_step), ("reference", ref_step), ("port", port_step)):
    fn(); t0 = time.perf_counter(); n = 5
    for _ in range(n): fn()
    dt = (time.perf_counter() - t0) / n
    print("%-10s %.3f s/step  %.3g edge-messages/s" % (name, dt, 4 * send.numel() / dt))
