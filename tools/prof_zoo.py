import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fedb200.control.trainer import LocalTrainer, TrainerConfig
from fedb200.utils import data as D
name = sys.argv[1]
ds = D.synthetic_cifar(128 * int(os.environ.get('ZOO_STEPS', '2')), 128, seed=0)
t = LocalTrainer(TrainerConfig(model=name, device="cuda:0", seed=0, use_cuda_graph=False), ds)
t.train(1, 0, 1)
torch.cuda.synchronize()
