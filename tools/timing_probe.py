import os, sys, numpy as np, torch
sys.path.insert(0, os.getcwd())
from metta_b200.batch import Batch
from metta_b200.builders import make_arena
from metta_b200.lowering import lower_game_config
from metta_b200.maps import arena_mapgen_pool, build_pool
N=int(sys.argv[1]) if len(sys.argv)>1 else 4096
FLUSH=(sys.argv[2]=='flush') if len(sys.argv)>2 else True
A,T=24,200
cfg=make_arena(A); low=lower_game_config(cfg.game,62,62)
pool=build_pool(low, arena_mapgen_pool(low))
b=Batch(low,pool,N)
dev='cuda'
obs=torch.zeros((N*A,T,3),dtype=torch.uint8,device=dev); term=torch.zeros(N*A,dtype=torch.uint8,device=dev); trunc=torch.zeros_like(term); rew=torch.zeros(N*A,dtype=torch.float32,device=dev)
act=torch.randint(0,5,(N*A,),dtype=torch.int32).to(dev)
b.set_buffers(obs.data_ptr(),term.data_ptr(),trunc.data_ptr(),rew.data_ptr(),act.data_ptr())
b.reset((np.arange(N)+42).astype(np.uint32),(np.arange(N)%256).astype(np.int32))
flush=torch.empty(256*1024*1024,dtype=torch.uint8,device=dev)
STEPS=int(sys.argv[3]) if len(sys.argv)>3 else 20
if len(sys.argv)>4 and sys.argv[4]=='stagger':  # the bench's steady state: envs at staggered episode phases
    g=np.arange(N,dtype=np.int64)
    b.set_early_reset((1+(g*6181)%9999).astype(np.uint32))
    for s in range(10000):
        act.copy_(torch.randint(0,5,(N*A,),dtype=torch.int32,device=dev)); b.step()
    torch.cuda.synchronize()
for s in range(STEPS):
    act.copy_(torch.randint(0,5,(N*A,),dtype=torch.int32))
    if FLUSH: flush.fill_(s)
    torch.cuda.synchronize()
    if s == STEPS-1: uses0 = b.read_assemblers()["uses_count"].astype(np.int64).sum(1)
    if s == STEPS-1: ord0 = b.read_env()["spare"][:, 0].astype(np.int64)
    if s == STEPS-1:
        prev_sp = b.read_buffer(16, np.uint32, (N, 16)).astype(np.float64); prev_sp[:,10] = 0; prev_tot = prev_sp.sum(1)
    b.step(); torch.cuda.synchronize()
uses = b.read_assemblers()["uses_count"].astype(np.int64).sum(1) - uses0
sp=b.read_buffer(16, np.uint32, (N, 16)).astype(np.float64)
names=['load: loads+fold','load: reset/stage','load: rest(prefetch,ctx)','decide: intents','decide: shuffle draws','decide: conflicts+apply','act parallel','act ordered','post','finish: rewards+records','', 'finish: write-back']
start=sp[:,10].copy(); sp[:,10]=0
tot=sp.sum(1)
print('N',N,'flush',FLUSH); print('per-env total cycles: mean %.0f p50 %.0f p90 %.0f p99 %.0f max %.0f'%(tot.mean(),np.percentile(tot,50),np.percentile(tot,90),np.percentile(tot,99),tot.max()))
print('envs with ordered work: %.1f%%' % (100.0*(sp[:,7] > 2000).mean()))
for i,n in enumerate(names):
    if not n: continue
    c=sp[:,i]; print('%-34s mean %8.0f p50 %8.0f p99 %8.0f max %8.0f'%(n,c.mean(),np.percentile(c,50),np.percentile(c,99),c.max()))

par, orde = sp[:,6], sp[:,7]
for k in range(0, 6):
    m = uses == k
    if m.sum(): print('envs with %d successful assembler uses in the step: %5d  act parallel mean %7.0f  ordered mean %7.0f  total mean %7.0f' % (k, m.sum(), par[m].mean(), orde[m].mean(), tot[m].mean()))
m = uses >= 6
if m.sum(): print('envs with >=6 uses: %d  act parallel mean %.0f total mean %.0f' % (m.sum(), par[m].mean(), tot[m].mean()))

su = sp[:,12:16]
m = uses == 1
if m.sum():
    print('one successful use: set-up+scans %.0f | checks %.0f | consume+give %.0f | record %.0f  (sum %.0f)' % (su[m,0].mean(), su[m,1].mean(), su[m,2].mean(), su[m,3].mean(), su[m].sum(1).mean()))

print('per-phase means by number of successful uses in the step (cycles):')
print('%-34s' % 'phase' + ''.join('%10s' % ('uses=%d' % k) for k in range(3)))
for i,n in enumerate(names):
    if not n: continue
    print('%-34s' % n + ''.join('%10.0f' % (sp[uses==k, i].mean() if (uses==k).any() else float('nan')) for k in range(3)))
m = sp[:,0] > 100000
if m.any():
    print('envs that started a new episode in this step: %d; load %.0f = accumulate %.0f | reset up to the assembler records %.0f | assembler records %.0f | count pass %.0f' % (
        m.sum(), sp[m,0].mean(), sp[m,12].mean(), sp[m,13].mean(), sp[m,14].mean(), sp[m,15].mean()))

rel=(start-start.min())%2**32
end=rel+tot/1.92
print('env start after the first one (ns): p50 %.0f p90 %.0f p99 %.0f max %.0f; started later than 10 us: %d envs' % (np.percentile(rel,50),np.percentile(rel,90),np.percentile(rel,99),rel.max(),(rel>10000).sum()))
print('env end (ns, start + cycles/1.92): p50 %.0f p90 %.0f p99 %.0f max %.0f' % (np.percentile(end,50),np.percentile(end,90),np.percentile(end,99),end.max()))
order=np.argsort(-end)[:16]
print('the envs that end last: env start end | load decide act-par act-ord post finish | uses')
for e in order:
    print('%5d %7.0f %7.0f | %6.0f %6.0f %6.0f %6.0f %6.0f %6.0f | %d' % (e, rel[e], end[e], sp[e,0:3].sum(), sp[e,3:6].sum(), sp[e,6], sp[e,7], sp[e,8], sp[e,9]+sp[e,11], uses[e]))

# is a slow env slow again in the next step? (what sorting envs into slices by their last chain length could buy)
ok = (prev_tot < 100000) & (tot < 100000)
r = np.corrcoef(prev_tot[ok], tot[ok])[0,1]
heavy_prev = prev_tot > np.percentile(prev_tot, 75); heavy_now = tot > np.percentile(tot, 75)
print('chain length, this step against the previous one: correlation %.3f; P(top quarter now | top quarter before) %.3f; P(top quarter now | not before) %.3f' % (r, heavy_now[heavy_prev].mean(), heavy_now[~heavy_prev].mean()))
very = tot > np.percentile(tot, 95)
print('of the slowest 5%% now, %.1f%% were in the top quarter before; the light three quarters (by the previous step) would end at p99 %.0f max %.0f cycles, all envs p99 %.0f max %.0f' % (100*heavy_prev[very].mean(), np.percentile(tot[~heavy_prev],99), tot[~heavy_prev & ok].max(), np.percentile(tot,99), tot[ok].max()))

# what would the kernel's end be if a phase were capped? (which tail is worth working on)
okk = sp[:,0] < 100000  # leave the env that started a new episode out
def kernel_end(adj_tot, first_wave_only=False):
    e = rel + adj_tot/1.92
    m = okk & ((rel < 10000) if first_wave_only else True)
    return e[m].max()
print('kernel end (ns), slowest env: now %.0f; ordered pass capped at 10k cycles %.0f; parallel pass capped at 15k %.0f; both %.0f; first wave only, now %.0f' % (
    kernel_end(tot), kernel_end(tot - np.maximum(0, sp[:,7]-10000)), kernel_end(tot - np.maximum(0, sp[:,6]-15000)),
    kernel_end(tot - np.maximum(0, sp[:,7]-10000) - np.maximum(0, sp[:,6]-15000)), kernel_end(tot, True)))

# what does one agent cost in the ordered pass?
n_ord = b.read_env()["spare"][:, 0].astype(np.int64) - ord0
m = (n_ord > 0) & okk
if m.any():
    A_ = np.vstack([n_ord[m], uses[m] * (sp[m, 7] > sp[m, 6]), np.ones(m.sum())]).T
    coef = np.linalg.lstsq(A_, sp[m, 7], rcond=None)[0]
    print('ordered pass: %d envs with ordered agents, %.1f agents each on average (max %d); cycles ~ %.0f per agent + %.0f per successful use there + %.0f' % (
        m.sum(), n_ord[m].mean(), n_ord[m].max(), coef[0], coef[1], coef[2]))
    for k in (1, 2, 3, 4, 6, 8, 12):
        mk = m & (n_ord == k)
        if mk.any(): print('  %2d ordered agents: %4d envs, ordered pass mean %6.0f p90 %6.0f cycles' % (k, mk.sum(), sp[mk, 7].mean(), np.percentile(sp[mk, 7], 90)))
