import sys
sys.path.insert(0,'/root/repo')
import numpy as np, importlib
pkg = importlib.import_module("orb-slam2-_b200")
w,h,nf=640,480,1000
img=pkg.synth.frame(w,h,3)
pimg=pkg.pinned_empty((h,w),np.uint8); pimg[:]=img
ex=pkg.ORBextractor(nf,1.2,8,20,7,max_width=w,max_height=h)
cap=nf+64
k,d,n=pkg.pinned_empty((1,cap),pkg.KP_DTYPE),pkg.pinned_empty((1,cap,32),np.uint8),np.zeros(1,np.int32)
for i in range(6):
    ex.extract_batch_into(pimg[None],k,d,n)
