import sys, torch
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
from common import O
from coolchic_b200 import ops
dev=torch.device('cuda:0')
SYN=["16-1-linear-relu","3-1-linear-none","3-3-residual-relu","3-3-residual-none"]
for (H,W) in [(32,48),(64,96),(512,768)]:
    arch=O.Arch(img_size=(H,W),dim_arm=8,n_hidden_layers_arm=2,layers_synthesis=SYN)
    p=O.init_params(arch,seed=1)
    target=O.synthetic_image(3,H,W).to(dev)
    spec=ops.ArchSpec(H,W,8,2,SYN,7,8,7,16.0)
    st=ops.FusedState(spec,1,dev)
    def run(two):
        st.load(p); st.arena[1:].zero_(); st.step.zero_(); st.philox_offset.zero_()
        st.lr.fill_(1e-2); st.T.fill_(0.3); st.sigma.fill_(2.0)
        for _ in range(3):
            st.train_step(target,1e-3,"softround","kumaraswamy",seed=5,two_streams=two)
        torch.cuda.synchronize()
        return st.out4.clone(), st.arena.clone()
    o_ref,a_ref=run(False)
    worst=0
    for it in range(40):
        o,a=run(True)
        d=(a-a_ref).abs().max().item(); worst=max(worst,d)
        if not torch.isfinite(a).all() or d>1e-3:
            print('  BAD it',it,'maxdiff',d,o.tolist(),o_ref.tolist())
    print(H,W,'worst maxdiff two-stream vs serial',worst, o_ref.tolist())
