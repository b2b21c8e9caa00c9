"""Which model of a hypothesis wins a RANSAC tie?  (round 2: pair 18 of rank 0's config-3 bench batch landed 0.14 deg /
1.4 deg from cv2 although keypoints, matches and the hypothesis stream were identical.)

CPU only.  Replays OpenCV's RANSAC loop (oracle/geometry.py: same RNG, same update rule) on the pair's matched points
three times - with cv2's OWN minimal solver (cv2.findEssentialMat on exactly five points returns all its models,
stacked, in its own order), with the kernel's solver (tests/hostcheck instantiation) and with the numpy restatement -
and prints the per-hypothesis inlier counts side by side.  Finding: the model SETS always agreed; the ORDER of the
models of one hypothesis did not, because cv2 parametrises E over the null-space basis its Jacobi SVD fills in by
Gram-Schmidt of fixed pseudo-random vectors (not a LAPACK basis), and the order in which cv2.solvePoly returns the
roots depends on that parametrisation.  With two models of ONE hypothesis tied at 492 inliers (a near-double root)
cv2 kept one and the engine the other.  Since the solver uses cv2's basis (geom_math.cuh, oracle.cv2_null_space) the
three columns printed here are identical.

    python tools/diag_model_order.py [pair_index]
"""
import sys, numpy as np, cv2
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
import _hostcheck as hc
from oracle import geometry as og, orb as oo
from oracle.reference_path import PathConfig, ReferencePath
from bench_support.synth import synth_pair, synth_K
W,H=1280,720
pi=int(sys.argv[1]) if len(sys.argv)>1 else 18
a,b=synth_pair(3,pi,W,H)[:2]; K=synth_K(W,H)
rp=ReferencePath(K,PathConfig(nfeatures=5000,nlevels=12,match_mode='ratio',ratio=0.75,max_matches=500,ransac_max_iters=2000),canonical=True)
def feats(img):
    orb=cv2.ORB_create(nfeatures=5000,scaleFactor=1.1,nlevels=12,fastThreshold=15,scoreType=cv2.ORB_HARRIS_SCORE)
    kps,desc=orb.detectAndCompute(img,None)
    lv=oo.level_geometry(W,H,oo.OrbParams(5000,1.1,12))
    return oo.canonicalize_cv2(kps,desc,lv)
r1,r2=feats(a),feats(b)
m=rp.match(r1.desc,r2.desc)
pts1=np.stack([r1.x[m[:,0]],r1.y[m[:,0]]],axis=1).astype(np.float32)
pts2=np.stack([r2.x[m[:,1]],r2.y[m[:,1]]],axis=1).astype(np.float32)
def cv_solver(x1,x2):
    # cv2's own minimal solver on normalised points (K = I)
    Es,_=cv2.findEssentialMat(x1.astype(np.float64),x2.astype(np.float64),np.eye(3),method=cv2.RANSAC,prob=0.999,threshold=1.0)
    return [] if Es is None else [Es[3*i:3*i+3] for i in range(Es.shape[0]//3)]
def k_solver(x1,x2):
    return [e.reshape(3,3) for e in hc.five_point(x1,x2)]
trc,trk=[],[]
Ec,mc=og.find_essential_ransac(pts1,pts2,K,max_iters=2000,solver=cv_solver,trace=trc)
Ek,mk=og.find_essential_ransac(pts1,pts2,K,max_iters=2000,solver=k_solver,trace=trk)
E_cv,m_cv=cv2.findEssentialMat(pts1,pts2,K,method=cv2.RANSAC,prob=0.999,threshold=1.0,maxIters=2000)
print('iters cv-solver',len(trc),'kernel-solver',len(trk))
print('E cvsolver vs cv2', min(np.abs(Ec-E_cv).max(),np.abs(Ec+E_cv).max()), 'E kernel vs cv2', min(np.abs(Ek-E_cv).max(),np.abs(Ek+E_cv).max()))
for (i,s,c1),(j,s2,c2) in zip(trc,trk):
    flag = '' if c1==c2 else ('  <-- order' if sorted(c1)==sorted(c2) else '  <-- DIFF')
    print(i,s,c1,c2,flag)
tro=[]
Eo,mo=og.find_essential_ransac(pts1,pts2,K,max_iters=2000,trace=tro)
print('oracle numpy solver:')
for (i,s,c1),(j,s2,c2) in zip(trc,tro):
    flag = '' if c1==c2 else ('  <-- order' if sorted(c1)==sorted(c2) else '  <-- DIFF')
    print(i,c1,c2,flag)
