#!/usr/bin/env python
"""The native decode + augment batch call alone (csrc/runtime/batch_pipeline.cpp): 256 JPEGs of 256 x 256 already in
memory -> 3 x 227 x 227 crops, mirrored, mean-subtracted, as fp32 or bf16, on 1 and 8 threads; and the augment stage alone
on decoded pixels.  Host CPU only."""
import io, time, numpy as np, sys, torch
from PIL import Image
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cxxnet_b200.io import native
rng=np.random.default_rng(0)
x=np.linspace(0,8,256)
blobs=[]
for k in range(64):
    img=(np.sin(x+k)[:,None]*np.cos(x*1.3)[None,:]*80+128)[...,None]+rng.normal(0,12,(256,256,3))
    img=np.clip(img,0,255).astype(np.uint8)
    buf=io.BytesIO(); Image.fromarray(img).save(buf,format='JPEG',quality=88); blobs.append(buf.getvalue())
items=[(blobs[i%64], None, 256, 256, int(rng.integers(0,29)), int(rng.integers(0,29)), bool(i&1), 1.0, 0.0) for i in range(256)]
mean=np.array([120.,115.,100.],dtype=np.float32)
for dt,name in [(torch.float32,'fp32'),(torch.bfloat16,'bf16')]:
    dst=torch.zeros(256,3,227,227,dtype=dt)
    for th in (1,8):
        native.decode_augment_batch(items,(3,227,227),1,mean,1.0,dst,dt==torch.bfloat16,th)
        t=time.perf_counter(); reps=3 if th==1 else 8
        for _ in range(reps): st=native.decode_augment_batch(items,(3,227,227),1,mean,1.0,dst,dt==torch.bfloat16,th)
        dtm=(time.perf_counter()-t)/reps
        print("%s threads %d: %.0f images/s (%.2f ms per image per thread)" % (name, th, 256/dtm, dtm/256*1e3*th))
# decoded-pixels path (no JPEG)
px=[np.ascontiguousarray(np.asarray(Image.open(io.BytesIO(b))).transpose(2,0,1).astype(np.float32)) for b in blobs]
items2=[(None, px[i%64], 256, 256, 3, 5, bool(i&1), 1.0, 0.0) for i in range(256)]
dst=torch.zeros(256,3,227,227)
native.decode_augment_batch(items2,(3,227,227),1,mean,1.0,dst,False,1)
t=time.perf_counter()
for _ in range(3): native.decode_augment_batch(items2,(3,227,227),1,mean,1.0,dst,False,1)
print("augment only, 1 thread: %.3f ms per image" % ((time.perf_counter()-t)/3/256*1e3))
