import sys, types, math, torch
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
from common import O
from test_gpu_train import make_encoder, manager
from coolchic_b200.enc.training.train import train, test as evaluate
dev=torch.device('cuda:0')
H,W=32,48
bad=0
for rep in range(60):
  for i in range(2):
    target = O.synthetic_image(10 + i, H, W).to(dev)
    fe = make_encoder(H, W, dev); mgr = manager()
    logs=[]
    train(fe, target, mgr, max_iterations=100, frequency_validation=10, patience=1000, seed=i, val_logs=logs)
    l = evaluate(fe, target, mgr)
    if not math.isfinite(float(l.loss)) or l.psnr_db < 15:
        bad+=1
        print(rep, i, float(l.loss), l.psnr_db, l.rate_latent_bpp, [round(float(x.loss),5) for x in logs])
        sd=fe.state_dict()
        print('   nonfinite params:', [k for k,v in sd.items() if not torch.isfinite(v).all()][:5])
print('bad', bad)
