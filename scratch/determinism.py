# is train() bitwise repeatable from the same initialisation and seed?
import sys, copy, types
sys.path.insert(0, ".")
import torch
from coolchic_b200.enc.component.coolchic import CoolChicEncoderParameter
from coolchic_b200.enc.component.frame import FrameEncoder
from coolchic_b200.enc.training.train import train, test
from oracle import float_ref as O
dev = torch.device("cuda:0")
for (H, W, syn, C, nh) in [(48, 64, ["16-1-linear-relu", "3-1-linear-none"], 8, 1),
                           (512, 768, ["48-1-linear-relu", "3-1-linear-none", "3-3-residual-relu", "3-3-residual-none"], 16, 2)]:
    target = O.synthetic_image(5, H, W).to(dev)
    torch.manual_seed(1)
    prm = CoolChicEncoderParameter(layers_synthesis=syn, n_ft_per_res=[1] * 7, dim_arm=C, n_hidden_layers_arm=nh,
                                   ups_k_size=8, ups_preconcat_k_size=7, encoder_gain=16)
    prm.set_image_size((H, W))
    base = FrameEncoder(prm)
    out = []
    for rep in range(3):
        fe = copy.deepcopy(base); fe.to_device(dev)
        mgr = types.SimpleNamespace(lmbda=1e-3, iterations_counter=0, total_training_time_sec=0.0)
        for ph in range(2):
            train(fe, target, mgr, start_lr=1e-2, cosine_scheduling_lr=False, max_iterations=60, patience=100000,
                  frequency_validation=60, quantizer_type="softround", quantizer_noise_type="kumaraswamy",
                  softround_temperature=(0.3, 0.3), noise_parameter=(2.0, 2.0), seed=1000 * ph + 2)
            out.append(float(test(fe, target, mgr).loss))
    print(H, W, out)
