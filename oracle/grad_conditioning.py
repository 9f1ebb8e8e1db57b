"""How well conditioned are the gradients the parity tests compare?  TEST INFRASTRUCTURE (build container only).

Runs the CPU oracle (pinned to the reference, oracle/make_golden_real.py) on the `real_wavlm_step` case twice — in
float32, the reference's arithmetic, and in float64 — and prints, per parameter, how far the float32 gradient is from
the float64 one.  That distance is the floor for ANY fp32 implementation: a tolerance below it tests rounding order,
not correctness.  Output of the run that set the bars of tests/test_parity_real_gpu.py:
profiles/r02_grad_conditioning_wavlm.txt.

    python oracle/grad_conditioning.py
"""
import sys, torch
sys.path[:0]=['/root/repo','/root/repo/tests','/root/repo/oracle']
import dphubert_oracle as O
from conftest import load_golden
import test_parity_real_gpu as T
T.DEV=torch.device('cpu')
g=load_golden('real_wavlm_step')
class Stop(Exception): pass
cap={}
def fake(teacher, student, *a, **k):
    cap['t']=teacher; cap['s']=student; raise Stop
T.NativeDistillStep=fake
try: T.build_real_step(g, torch.float32)
except Stop: pass
teacher, student = cap['t'], cap['s']
cfg=T.CFG['wavlm'](); scfg=dict(cfg); scfg.update(extractor_prune_conv_channels=True, encoder_prune_attention_heads=True, encoder_prune_feed_forward_intermediate=True)
wave=T.make_wave(g['B'],g['T0'],g['wave_seed'],g['wave_sum']); lengths=torch.tensor([g['T0']]*g['B'])
D=768
res={}
for dt in (torch.float32, torch.float64):
    t_sd={k:v.detach().to(dt) if v.is_floating_point() else v for k,v in teacher.state_dict().items()}
    s_sd={k:(v.detach().to(dt).clone().requires_grad_("dummy" not in k) if v.is_floating_point() else v) for k,v in student.state_dict().items()}
    uniq={}; heads=[]
    for hid in g['head_ids']:
        if hid not in uniq: uniq[hid]=(torch.eye(D,dtype=dt).requires_grad_(True), torch.zeros(D,dtype=dt,requires_grad=True))
        heads.append(uniq[hid])
    lam=[torch.tensor(0.3,dtype=dt,requires_grad=True), torch.tensor(0.2,dtype=dt,requires_grad=True)]
    noise={k:v.to(dt) for k,v in g['noise'].items()}
    loss,_=O.distill_step_loss(t_sd,cfg,s_sd,scfg,heads,g['layers'],wave.to(dt),lengths,noise,lam,0.75,g['original_num_params'],(0.0,1.0,1.0),'raw','layer2layer')
    loss.backward()
    res[dt]={k:v.grad.double() for k,v in s_sd.items() if torch.is_tensor(v) and v.is_floating_point() and v.grad is not None}
worst=[]
for k in res[torch.float64]:
    a,b=res[torch.float32][k],res[torch.float64][k]
    if b.abs().max()<1e-9: continue
    worst.append(((a-b).norm().item()/b.norm().item(), (a-b).abs().max().item()/b.abs().max().item(), k))
worst.sort(reverse=True)
for w in worst[:12]: print("%.2e %.2e %s"%w)
print("median l2 %.2e"%sorted(w[0] for w in worst)[len(worst)//2])
