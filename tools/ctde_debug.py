import numpy as np, torch, sys, ctypes as C
sys.path.insert(0, '.')
from oracle import ctde as oc
from tianshou_marl_b200 import _lib
from tianshou_marl_b200.algorithm.ctde import CentralizedCritic, DecentralizedActor, GlobalStateConstructor, load_linear3
g = np.load('tests/golden/ctde.npz'); pre = 'c0_'
c = {k[len(pre):]: g[k] for k in g.files if k.startswith(pre)}
A, O, H, nA = int(c['A']), int(c['O']), int(c['H']), int(c['nA'])
obs = torch.tensor(c['obs'], device='cuda'); obs_next = torch.tensor(c['obs_next'], device='cuda')
gs = GlobalStateConstructor('concatenate', n_agents=A)
cin, cinn = gs.build_rows(obs), gs.build_rows(obs_next)
critic = load_linear3(CentralizedCritic(A * O, A, H), {k[len('init_critic_'):]: v for k, v in c.items() if k.startswith('init_critic_')})
mk = lambda n: oc.Mlp3(*(c[f'init_{n}_fc{i}.{w}'] for i in (1, 2, 3) for w in ('weight', 'bias')))
oc_c = mk('critic')
with torch.no_grad():
    ref_v = oc_c(cin.cpu()); ref_vn = oc_c(cinn.cpu())
n = obs.shape[0]
v = torch.empty(n, A, device='cuda'); vn = torch.empty(n, A, device='cuda')
critic.forward(cin, mode='raw', out={'out': v}); critic.forward(cinn, mode='raw', out={'out': vn})
print('critic fwd diff', (v.cpu() - ref_v).abs().max().item(), (vn.cpu() - ref_vn).abs().max().item())
rew = torch.tensor(c['rew'], device='cuda'); term = torch.tensor(c['term'], device='cuda').to(torch.uint8)
td = torch.empty(n, device='cuda'); adv = torch.empty(n, device='cuda')
L = _lib.lib(); P = _lib.ptr; S = _lib.stream_ptr()
_lib.check(L.tsm_td_target(P(v), P(vn), A, P(rew), P(term), float(c['gamma']), n, P(td), P(adv), None, S))
ref_td = torch.from_numpy(c['rew']) + float(c['gamma']) * ref_vn.mean(1) * (~torch.from_numpy(c['term'])).float()
print('td diff', (td.cpu() - ref_td).abs().max().item(), 'mse ref', ((ref_v.mean(1) - ref_td) ** 2).mean().item())
pb = _lib.TsmPpoBatch(obs=P(cin), act=None, adv=None, ret=P(td), logp_old=None, v_old=None, indices=None, size=1, n_rows=n, start=0, len=1, perm_key=0, perm_n=0, records=None)
cfg = _lib.TsmPpoCfg(eps_clip=0.2, dual_clip=0.0, vf_coef=1.0, ent_coef=0.0, value_clip=0, adv_norm=0, plain_pg=0, value_heads=A)
sums = torch.zeros(4, dtype=torch.float64, device='cuda')
_lib.check(L.tsm_ppo_minibatch(critic.desc, C.byref(pb), C.byref(cfg), 0, n, None, P(sums), S))
print('sums', sums.cpu().numpy() / n)
