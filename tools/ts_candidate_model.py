"""Numpy model of the lazy Gumbel-max sampler candidate counts under the hierarchical noise (64-code super-groups / groups of 8):
how many super-groups, groups and codes per row pass the bound tests with a scout-only threshold and with the final one.
    python tools/ts_candidate_model.py > profiles/r02_candidate_model.txt
"""
import numpy as np
rng=np.random.default_rng(0)
D,K,T=24,8192,0.03
R=2000
z=rng.standard_normal((R,D)); z/=np.linalg.norm(z,axis=1,keepdims=True)
e=rng.uniform(-1,1,(K,D)); e/=np.linalg.norm(e,axis=1,keepdims=True)
s=z@e.T
eps=4.2e-3
# hierarchical noise in E-space
NS=K//64
XS=rng.exponential(size=(R,NS)); ES=XS/64
hs13=np.floor((1-np.exp(-XS))*8192)  # w = 1-exp(-X) uniform; head = floor(w*2^13)
ESlb=hs13/2**19
JS=rng.integers(0,8,(R,NS))
Y=rng.exponential(size=(R,NS,8)); hg14=np.floor((1-np.exp(-Y))*16384)
EG=ES[:,:,None]+Y/8
for r in range(R):
    pass
mask=np.arange(8)[None,None,:]==JS[:,:,None]
EG=np.where(mask,ES[:,:,None],EG)
EGlb=ESlb[:,:,None]+np.where(mask,0,hg14/2**17)
J=rng.integers(0,8,(R,NS,8))
Ep=EG[...,None]+rng.exponential(size=(R,NS,8,8))
pm=np.arange(8)[None,None,None,:]==J[...,None]
Ep=np.where(pm,EG[...,None],Ep)
g=-np.log(Ep).reshape(R,K)
v=s/T+g
vstar=v.max(1)
def count(scout_tiles, label):
    nsc=scout_tiles*256
    # threshold: best approx code among scout codes, its true value (lower bound w/ eps)
    k1=s[:,:nsc].argmax(1)
    thr=(s[np.arange(R),k1]-eps)/T+g[np.arange(R),k1]
    mS=s.reshape(R,NS,64).max(2)
    with np.errstate(divide='ignore'):
        ubS=(mS+eps)/T-np.log(ESlb)
    passS=ubS>=thr[:,None]
    mG=s.reshape(R,NS,8,8).max(3)
    with np.errstate(divide='ignore'):
        ubG=(mG+eps)/T-np.log(EGlb)
    passG=(ubG>=thr[:,None,None])&passS[:,:,None]
    # code-level with group max only
    gg=g.reshape(R,NS,8,8)
    passC=((mG[...,None]+eps)/T+gg>=thr[:,None,None,None])&passG[...,None]
    # with final threshold
    passS2=ubS>=vstar[:,None]; passG2=(ubG>=vstar[:,None,None])&passS2[:,:,None]
    passC2=((mG[...,None]+eps)/T+gg>=vstar[:,None,None,None])&passG2[...,None]
    sc=s.reshape(R,NS,8,8)
    passC3=((sc+eps)/T+gg>=vstar[:,None,None,None])&passG2[...,None]
    print(f"{label}: scout thr: supers/row {passS.sum()/R:.1f} groups/row {passG.sum()/R:.1f} codes/row {passC.sum()/R:.1f} | final thr: supers {passS2.sum()/R:.1f} groups {passG2.sum()/R:.1f} codes(mg) {passC2.sum()/R:.1f} codes(percode) {passC3.sum()/R:.1f}  gap {np.mean(vstar-thr):.2f}")
for st in (1,2,4,8,16,32): count(st,f"scout {st}")
