import sys, numpy as np
sys.path.insert(0,'.')
from oracle import jet as ojet
from radiosim_b200.jets.jet import render_batch
from radiosim_b200.jets.packer import pack_jets
for size,n,seed in [(128,12,1),(200,12,2),(256,12,3),(256,2,11)]:
    np.random.seed(seed); tr=[]
    ox,_=ojet.create_jet((n,3,size,size),(3,10),"normalize","clean",trace=tr)
    np.random.seed(seed); b=pack_jets(n,size,(3,10),"normalize")
    img,margin,peak,_=render_batch(b,True,f64=False)
    for i in range(n):
        e=np.abs(img[i]-ox[i,0]); r,c=np.unravel_index(e.argmax(),e.shape)
        print(size,i,'zoom',int(b.zoom_pad[i]),'m',margin[i],tr[i]['margin'],'err %.2e at'%e.max(),r,c,'peak %.6f %.6f'%(peak[i],tr[i]['jet_max']), 'rowerr', np.where(e.max(axis=1)>1e-4)[0][:6], 'colerr', np.where(e.max(axis=0)>1e-4)[0][:6])
