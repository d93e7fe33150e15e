"""Config 3 / the reference's only runnable entry point (cifar_example/sgd-torchad_nn-cifar.lua):
CIFAR-10-shaped CNN, automatic data parallelism through parallelize().

    examples/launch.sh 8 examples/cifar_dp.py -usegpu -batchSize 1 -iterations 1 --miniBatch 4096
    python examples/cifar_dp.py                       # CPU, 1 rank, reference defaults (tiny)

Same 8 flags and defaults as the Lua script (:29-43), same results-file fields (:277-299), the
confusion matrix of :266-275.  --dataDir points at the reference's dataset files
(cifar-10-batches-t7/*.t7, :133-169); without it (there is no network on this box) the dataset is
synthetic CIFAR-shaped data (class-dependent mean, learnable).  -data selects 32000/10000 instead of
4000/2000 samples as in the reference (:124-130).  Preprocessing (rgb2yuv, U/V standardisation with
the training shard's statistics, Y contrast normalisation, :182-216) runs after parallelize() as in
the reference — through the native kernels of csrc/preprocess.cu on a GPU, the tensor implementation
on the CPU; --preprocess none skips it.  --miniBatch > 1 is the B200 extension: the reference trains
one sample per step.
"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tadnn as automation  # noqa: E402
import tadnn.mpi as mpi  # noqa: E402
from tadnn import nn  # noqa: E402
from tadnn.models import CifarCNN  # noqa: E402
from tadnn.utils import config as cfg  # noqa: E402
from tadnn.utils import preprocess as prep  # noqa: E402
from tadnn.utils.cifar import load_cifar10_t7  # noqa: E402
from tadnn.utils.confusion import ConfusionMatrix  # noqa: E402
from tadnn.utils.logging import log, write_results_file  # noqa: E402
from tadnn.utils.timing import Tic  # noqa: E402

CLASSES = ["airplane", "automobile", "bird", "cat", "deer", "dog", "frog", "horse", "ship", "truck"]


def synthetic_cifar(n, seed):
    g = torch.Generator().manual_seed(seed)
    labels = torch.randint(0, 10, (n,), generator=g)
    protos = torch.randn(10, 3, 32, 32, generator=torch.Generator().manual_seed(1234))
    x = 0.7 * torch.randn(n, 3, 32, 32, generator=g) + protos[labels]
    return x, labels


def main():
    ap = cfg.add_reference_flags(argparse.ArgumentParser(description="CIFAR Training"))
    opt = cfg.from_args(ap.parse_args())
    tic = Tic()
    torch.manual_seed(opt.seed)          # :48 — same seed on every rank
    torch.set_num_threads(opt.threads)   # :51
    use_gpu = opt.usegpu and torch.cuda.is_available()
    h = mpi.start(use_gpu)
    dev = h.device()
    model = CifarCNN(seed=opt.seed)
    if use_gpu:
        model = model.to(dev)            # :239-245 (moved before parallelize so the flat buffer is symmetric)
    if opt.dataDir:
        # :133-169 — the Torch7 batch files the reference downloads; here they must already be on disk
        train_x, train_y, test_x, test_y = load_cifar10_t7(opt.dataDir, opt.trsize, opt.tesize)
    else:
        train_x, train_y = synthetic_cifar(opt.trsize, opt.seed)
        test_x, test_y = synthetic_cifar(opt.tesize, opt.seed + 1)
    if use_gpu:
        train_x, test_x = train_x.bfloat16(), test_x.bfloat16()
    # :172 — the one call
    train_x, train_y, trsize = automation.parallelize(train_x, train_y, model, opt.trsize, h, None, opt.batchSize)
    if opt.preprocess == "yuv":
        # :182-216 — statistics come from this rank's training shard and are applied to the test set
        if use_gpu:
            tr, te = prep.preprocess_cuda(train_x.to(dev).float().contiguous(), test_x.to(dev).float().contiguous(),
                                          out_dtype=torch.bfloat16)
            train_x, test_x = tr.cpu(), te.cpu()
        else:
            train_x, test_x = prep.preprocess(train_x.float().contiguous(), test_x.float())
    t1 = tic.toc()
    tic.tic()
    if opt.miniBatch > 1:
        trainer = automation.BatchTrainer(model, learningRate=opt.learningRate, weightDecay=opt.weightDecay)
        mb = min(opt.miniBatch, trsize)
        steps = max(1, trsize // mb) * opt.iterations
        xs = train_x.pin_memory() if use_gpu else train_x
        ys = train_y.pin_memory() if use_gpu else train_y
        last = None
        for last in trainer.fit(xs, ys, mb, steps):
            pass
        log(f"trained {steps} steps of {mb} samples, last loss {last:.4f}")
    else:
        # the reference's per-sample loop (datamodule.lua:117-184), CrossEntropyCriterion on log-probs
        # On a GPU the model itself goes to the trainer: StochasticGradient then runs every sample as one replay of
        # a captured CUDA graph of the fused tadnn step (CifarCNN.native_sample_step).  On the CPU the same network
        # as a Sequential of stock layers.
        seq = model if use_gpu else model.to_sequential()
        seq._tadnn_sync = getattr(model, "_tadnn_sync", None)
        trainer = nn.StochasticGradient(seq, nn.CrossEntropyCriterion())
        trainer.learningRate, trainer.maxIteration = opt.learningRate, opt.iterations
        trainer.train((train_x.to(dev).float(), train_y.to(dev)))
    if use_gpu:
        torch.cuda.synchronize()
    t2 = tic.toc()
    # :266-275 — every rank evaluates the full test set in the reference; here rank 0 reports
    confusion = ConfusionMatrix(CLASSES)
    model.eval()
    with torch.no_grad():
        for i in range(0, opt.tesize, 1000):
            xb = test_x[i:i + 1000].to(dev)
            confusion.add(model(xb).float().cpu(), test_y[i:i + 1000])
    if h.rank() == 0:
        print(confusion)
        path = write_results_file(opt.results_dir, train_size=trsize, test_size=opt.tesize, batch_size=opt.batchSize,
                                  accuracy=confusion.totalValid * 100, learning_rate=opt.learningRate,
                                  threads=opt.threads, preprocess_s=t1, train_s=t2)
        log(f"results written to {path}; {trsize * max(1, opt.iterations) / t2:.1f} samples/s/rank wall-clock")
    h.stop()


if __name__ == "__main__":
    main()
