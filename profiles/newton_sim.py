#!/usr/bin/env python
"""CPU simulation (numpy Newton, statistics only) of the lockstep waste of the curve solver and of what tail compaction / a sorted work
list would save; the numbers quoted in DESIGN.md section 8. python profiles/newton_sim.py"""
import json, os, sys, numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from optimo_b200 import synthetic
from optimo_b200.core import Core
def sim(scene, B, ntri_per_group=None):
    c=Core(scene); x0=c.GetDescriptor()
    X=synthetic.perturbed_candidates(x0,B)
    # channels: list of cp index ranges
    chans=[]; cp=0
    for j in synthetic.iter_joints(scene['joints']):
        for k in ('rot_x','rot_y','rot_z'):
            n=len(j[k]['curve'])
            if n>0 and j is not scene['joints']: chans.append((cp,n))
            cp+=n
    lo,hi=c.window
    t=[]; tt=lo
    while tt<=hi: t.append(tt); tt+=0.2
    t=np.array(t)
    its=np.zeros((B,len(chans),len(t)),dtype=np.int32)
    for b in range(B):
        x=X[b].reshape(-1,5)
        for ci,(c0,n) in enumerate(chans):
            P=x[c0:c0+n]; T=P[:,0]
            m=np.searchsorted(T,t,side='right')
            ok=(m>0)&(m<n)
            sp=np.clip(m-1,0,n-1); ep=np.clip(m,0,n-1)
            p0=T[sp]; p1=T[sp]+P[sp,4]; p2=T[ep]+P[ep,3]; p3=T[ep]
            a=-p0+3*p1-3*p2+p3; bb=3*(p0-2*p1+p2); cc=-3*p0+3*p1; d=p0-t
            u=np.full(t.shape,0.5); k=np.zeros(t.shape,dtype=np.int32)
            act=ok.copy()
            for it in range(50):
                f=u*(u*(a*u+bb)+cc)+d; fp=u*(3*a*u+2*bb)+cc
                act=act&~(np.abs(f)<1e-9)
                if not act.any(): break
                u=np.where(act,u-f/np.where(fp==0,1,fp),u); k+=act
            its[b,ci]=k
    return its
def cost(its, thresh):
    B,C,N=its.shape
    # warps of 32 consecutive samples (tiles of 124 ignored), triples of channels
    tot_lock=0; tot_new=0; tot_ideal=0
    for b in range(B):
        for w0 in range(0,N-31,32):
            for c0 in range(0,C,3):
                k=its[b,c0:c0+3,w0:w0+32]   # [<=3,32]
                K=k.shape[0]
                tot_ideal+=k.sum()/32.0
                tot_lock+=k.max()*K
                # new: lockstep rounds while active pairs > thresh, then tail
                r=0
                while True:
                    act=(k>r).sum()
                    if act<=thresh: break
                    r+=1
                rem=np.clip(k-r,0,None).sum(axis=0)  # per lane remaining
                tot_new+=r*K+rem.max()*1.15
    return tot_lock,tot_new,tot_ideal
scene=synthetic.stick_240(json.load(open(ROOT + '/tests/golden/scenes/stick-figure.json')))
its=sim(scene,8)
print('stick240 mean iters (in-range)',its[its>0].mean(), 'max',its.max())
for th in (0,16,24,32,40,48,64,96):
    l,n,i=cost(its,th); print(th, 'lock',l,'new',round(n),'ideal',round(i), 'ratio',round(n/l,3))
scene=synthetic.tree_scene(64,200)
its=sim(scene,2)
print('tree mean iters',its[its>0].mean(),'max',its.max())
for th in (0,16,32,48,64,96):
    l,n,i=cost(its,th); print(th,'lock',l,'new',round(n),'ideal',round(i),'ratio',round(n/l,3))

print("---- sorted lockstep simulation ----")
def cost_sorted(its, its_proxy, G):
    """its [B,C,N]; proxy [C,N] predicted cost; groups of G channels; tiles of 124 owned + 4 halo (128-window)."""
    B,C,N=its.shape
    tot_lock=tot_sorted=tot_ideal=0
    for b in range(B):
        for tile0 in range(0,N,124):
            s0=max(tile0-2,0); s1=min(s0+128,N)
            if s1-s0<128: continue
            for g0 in range(0,C,G):
                k=its[b,g0:g0+G,s0:s1]          # [g,128]
                g=k.shape[0]
                ntri=(g+2)//3
                # current: per warp, triples of channels
                for w in range(4):
                    for t in range(ntri):
                        kk=k[3*t:3*t+3, 32*w:32*w+32]
                        tot_lock+=kk.max()*3
                tot_ideal+=k.sum()/32.0
                # sorted: order items by proxy descending, chunks of 96
                pr=its_proxy[g0:g0+G,s0:s1].reshape(-1)
                order=np.argsort(-pr,kind='stable')
                kf=k.reshape(-1)[order]
                pad=(-len(kf))%96
                kf=np.concatenate([kf,np.zeros(pad,dtype=kf.dtype)])
                for c0 in range(0,len(kf),96):
                    tot_sorted+=kf[c0:c0+96].max()*3
    return tot_lock,tot_sorted,tot_ideal
scene=synthetic.stick_240(json.load(open(ROOT + '/tests/golden/scenes/stick-figure.json')))
its=sim(scene,8)
# proxy A: iteration counts of the unperturbed x0 (candidate-independent)
c=Core(scene); x0=c.GetDescriptor()
import copy
def sim_x(scene,x):
    # reuse sim with B=1 by monkeypatching perturbed_candidates
    old=synthetic.perturbed_candidates
    synthetic.perturbed_candidates=lambda x0,B,**k: np.tile(x,(1,1))
    r=sim(scene,1); synthetic.perturbed_candidates=old; return r[0]
proxyA=sim_x(scene,x0)
print('stick240 G=6 proxy=x0 iters', cost_sorted(its,proxyA,6))
# proxy B: mean over a few candidates
proxyB=its[:2].mean(axis=0)
print('stick240 G=6 proxy=mean of cand 0,1', cost_sorted(its[2:],proxyB,6))
print('stick240 oracle proxy (own counts)', cost_sorted(its[:1],its[0],6))
scene=synthetic.tree_scene(64,200)
its=sim(scene,3)
c=Core(scene); x0=c.GetDescriptor()
proxyA=sim_x(scene,x0)
print('tree G=12 proxy=x0', cost_sorted(its,proxyA,12))
print('tree G=12 proxy=cand0', cost_sorted(its[1:],its[0],12))

print("---- sample permutation (all channels of a group keep their lane; lanes take samples sorted by predicted cost) ----")
def cost_sample_perm(its, proxy, G):
    B,C,N=its.shape
    tot_lock=tot_perm=0
    for b in range(B):
        for tile0 in range(0,N,124):
            s0=max(tile0-2,0); s1=min(s0+128,N)
            if s1-s0<128: continue
            for g0 in range(0,C,G):
                k=its[b,g0:g0+G,s0:s1]
                g=k.shape[0]; ntri=(g+2)//3
                for w in range(4):
                    for t in range(ntri):
                        tot_lock+=k[3*t:3*t+3,32*w:32*w+32].max()*3
                pr=proxy[g0:g0+G,s0:s1].sum(axis=0)           # predicted cost of a sample = sum over the group's channels
                order=np.argsort(-pr,kind='stable')
                kp=k[:,order]
                for w in range(4):
                    for t in range(ntri):
                        tot_perm+=kp[3*t:3*t+3,32*w:32*w+32].max()*3
    return tot_lock,tot_perm
scene=synthetic.tree_scene(64,200)
its=sim(scene,3)
c=Core(scene); x0=c.GetDescriptor()
proxyA=sim_x(scene,x0)
print('tree G=12 sample-perm proxy=x0', cost_sample_perm(its,proxyA,12))
scene=synthetic.stick_240(json.load(open(ROOT + '/tests/golden/scenes/stick-figure.json')))
its=sim(scene,8); c=Core(scene); x0=c.GetDescriptor(); proxyA=sim_x(scene,x0)
print('stick240 G=6 sample-perm proxy=x0', cost_sample_perm(its,proxyA,6))
