"""worker for test_gpu_multi.py: run under torchrun, one rank per GPU.  argv[1] = updater."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
UPD = sys.argv[1] if len(sys.argv) > 1 else "sgd"
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch, torch.distributed as dist
rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
from util import make_trainer, make_batch, rel_err, copy_weights
from cxxnet_b200.models import lenet_conf
from cxxnet_b200.trainer import NetTrainer
extra = "wmat:wd = 0.001\nupdater = " + UPD + "\n"
_lenet = lenet_conf
def lenet_conf(**kw):   # dropout off: masks are a function of the element index, i.e. of the batch split
    return _lenet(**kw).replace("threshold = 0.5", "threshold = 0.0")
MODE = sys.argv[2] if len(sys.argv) > 2 else "lenet"
COLLECTIVE = MODE == "collective"       # torch.distributed all-reduce + local update instead of the fused kernels
if COLLECTIVE:
    os.environ["CXN_EXCHANGE"] = "collective"
    MODE = "lenet"
GRAPH = len(sys.argv) > 3 and sys.argv[3] == "graph"     # multi-GPU trainer replays captured steps
if MODE == "gather":
    # fully connected net whose first layer can be row-sharded (fullc_gather = 1): nin = 1024
    def lenet_conf(batch_size=64, dev="gpu", extra=""):
        return """
netconfig=start
layer[+1] = fullc:cv1
  nhidden = 96
  fullc_gather = 1
layer[+1] = relu
layer[+1] = fullc:fc1
  nhidden = 40
layer[+1] = sigmoid
layer[+1] = fullc:fc2
  nhidden = 10
layer[+0] = softmax
netconfig=end
input_shape = 1,1,1024
batch_size = %d
dev = %s
random_type = xavier
eta = 0.1
momentum = 0.9
metric = error
%s
""" % (batch_size, dev, extra)
    SHAPE = (64, 1, 1, 1024)
elif MODE == "shard":
    # three fully connected layers sharded by output columns (fullc_shard = 1, parallel/tensor_parallel.py):
    # row gather -> column gathers -> all-to-all forward, all-to-all -> pull reductions backward
    def lenet_conf(batch_size=64, dev="gpu", extra=""):
        return """
netconfig=start
layer[+1] = fullc:cv1
  nhidden = 128
layer[+1] = relu
layer[+0] = dropout
  threshold = 0.0
layer[+1] = fullc:fc1
  nhidden = 64
layer[+1] = sigmoid
layer[+1] = fullc:fc2
  nhidden = 32
layer[+0] = softmax
netconfig=end
input_shape = 1,1,1024
batch_size = %d
dev = %s
random_type = xavier
eta = 0.1
momentum = 0.9
metric = error
fullc_shard = 1
%s
""" % (batch_size, dev, extra)
    SHAPE = (64, 1, 1, 1024)
else:
    SHAPE = (64, 1, 28, 28)
tr = make_trainer(lenet_conf(batch_size=64, dev="gpu:" + ",".join(str(i) for i in range(world)),
                             extra=extra + ("cuda_graph = 1\n" if GRAPH else "")))
if MODE == "gather":
    assert len(tr.net.fc_shards) == 1, "first layer should be sharded"
if MODE == "shard":
    assert tr.net.tp is not None and len(tr.net.tp.chains) == 1 and len(tr.net.tp.chains[0]) == 3, "three sharded layers"
    assert not tr.net.exchange.buckets, "no gradient buckets: every weighted layer is column-sharded"
assert tr.net.exchange is not None and tr.world_size == world
assert (type(tr.net.exchange).__name__ == "CollectiveExchange") == COLLECTIVE
os.environ.pop("CXN_EXCHANGE", None)
assert tr.net.exchange.check_replicas() == 0.0
# single-GPU twin: same weights, whole batch on one device, no exchange
single = make_trainer(lenet_conf(batch_size=64, dev="gpu:%d" % rank, extra=extra), param_server="NONE")
assert single.world_size == 1 and single.net.exchange is None
copy_weights(tr, single)
# phase A: gradients only (no update): sum over ranks must equal the single-device gradient
b = make_batch(SHAPE, seed=99)
for t in (tr, single):
    lo, hi = t._slice(b.batch_size)
    t.net.label_info.fields = t._label_fields(b.label, lo, hi, t.device)
    t.net.accumulate_grads = False
    t.net.set_input(b.data[lo:hi], [])
    t.net.forward(True)
    t.net.backprop(False, False, 0)
torch.cuda.synchronize()
for i, c in enumerate(tr.net.connections):
    for tag, w, g in c.layer.weights():
        gm = tr.net.get_grad(i, tag).clone()
        if not (i in getattr(tr.net, "fc_shards", {}) and tag == "bias"):
            dist.all_reduce(gm)      # (a sharded layer's bias gradient is already the full-batch sum on every rank)
        gs = single.net.get_grad(i, tag)
        e = rel_err(gs, gm, floor=1e-7)
        if rank == 0:
            print("  grad[%d:%s] rel err %.4g |single| %.4g |multi| %.4g" % (i, tag, e, float(gs.abs().sum()), float(gm.abs().sum())))
        assert e < 2e-2, (i, tag, e)
# clear accumulated gradients again
tr.net.arena.g.zero_(); single.net.arena.g.zero_()
torch.cuda.synchronize(); dist.barrier()
for s in range(9 if GRAPH else 4):
    b = make_batch(SHAPE, seed=s)
    tr.update(b)
    single.update(b)
torch.cuda.synchronize()
if GRAPH:
    assert len(tr._graphs) == 2, "expected one captured graph per staging slot, got %d" % len(tr._graphs)
assert tr.net.exchange.check_replicas() == 0.0, "replicas diverged"
worst = 0.0
for name in ("cv1", "fc1", "fc2"):
    for tag in ("wmat", "bias"):
        a, _ = tr.get_weight(name, tag)
        b_, _ = single.get_weight(name, tag)
        e = rel_err(torch.from_numpy(b_), torch.from_numpy(a))
        if rank == 0:
            print("  %s.%s rel err %.4g  |single| %.4g |multi| %.4g" % (name, tag, e, abs(b_).sum(), abs(a).sum()))
        worst = max(worst, e)
print("RANK", rank, "worst rel err", worst)
assert worst < 3e-2, worst
dist.barrier()
dist.destroy_process_group()
print("OK", rank)
