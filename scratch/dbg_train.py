import sys, types, math, torch
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
from common import O
from test_gpu_train import make_encoder, manager
from coolchic_b200.enc.training.train import train, test as evaluate
dev=torch.device('cuda:0')
H,W=32,48
for rep in range(3):
  for i in range(2):
    target = O.synthetic_image(10 + i, H, W).to(dev)
    fe = make_encoder(H, W, dev); mgr = manager()
    logs=[]
    train(fe, target, mgr, max_iterations=100, frequency_validation=50, patience=1000, seed=i, val_logs=logs)
    l = evaluate(fe, target, mgr)
    print(rep, i, float(l.loss), l.psnr_db, l.rate_latent_bpp, [float(x.loss) for x in logs])
