"""Command line of the split-NN runs: same flags, defaults, validation and per-mode call order as
`/root/reference/split_nn.py` (argparse :152-176, `example()` :34-146).

The reference spawns world_size processes joined by torch.distributed.rpc (rank 0 = Bob, the
others idle RPC servers hosting one Alice each).  Here one process drives every entity: Bob and the
Alices are plain objects whose methods enqueue CUDA work; with several GPUs visible Bob sits on
cuda:0 and Alice r on cuda:(1 + (r-1) mod (n_gpus-1)) (`place_entities`; every entity logs its
device), the producing kernel storing activations / cut-layer gradients straight into the consuming
GPU's memory over NVLink with events ordering the streams.  `--world_size` keeps its meaning:
1 server + (world_size-1) clients.  (The one-process-per-GPU form that bench.py times is spmd.py.)

    python -m splitlearning_b200.split_nn --vanilla --world_size 5 --epochs 1 --iterations 1
"""
import argparse
import os

import torch


def get_alice_and_bob(args):
    if args.vanilla:
        from .data_entities_vanilla import alice, bob
    elif args.concat:
        from .data_entities_sisa_concat import alice, bob
    elif args.sisa:
        from .data_entities_vanilla_sisa import alice, bob
    else:
        from .data_entities import alice, bob
    return alice, bob


def build_parser():
    parser = argparse.ArgumentParser(description='Split Learning Initialization')
    parser.add_argument('--world_size', type=int, default=3, help='The world size which is equal to 1 server + (world size - 1) clients')
    parser.add_argument('--epochs', type=int, default=1, help='The number of epochs to run on the client training each iteration')
    parser.add_argument('--iterations', type=int, default=5, help='The number of iterations to communication between clients and server')
    parser.add_argument('--batch_size', type=int, default=16, help='The batch size during the epoch training')
    parser.add_argument('--partition_alpha', type=float, default=0.5, help='Number to describe the uniformity during sampling (heterogenous data generation for LDA)')
    parser.add_argument('--datapath', type=str, default="data/mnist_flat", help='folder path to all the local datasets')
    parser.add_argument('--lr', type=float, default=0.001, help='Learning rate of local client (SGD)')
    parser.add_argument('--server_epochs', type=int, default=3, help='The number of epochs to run on the server training each iteration')
    parser.add_argument('--vanilla', action='store_true', help='The trigger to switch over vanilla or U-shape split learning')
    parser.add_argument('--sisa', action='store_true', help='The trigger to switch over sisa and non-sisa')
    parser.add_argument('--concat', action='store_true', help='The trigger to use a concat layer')
    parser.add_argument('--control', action='store_true', help='The trigger to run for the control group or not')
    # additions (absent from the reference; defaults keep its behaviour)
    parser.add_argument('--synthetic_samples', type=int, default=70000, help='size of the synthetic MNIST-shaped set used when no mnist_784.npz is under --datapath')
    parser.add_argument('--seed', type=int, default=None, help='seed numpy/random (partition) and give every entity its own torch stream (Bob 100, Alice r 200+r)')
    parser.add_argument('--n_gpus', type=int, default=None, help='use at most this many of the visible GPUs (Bob on cuda:0, Alices on the others); default: all')
    parser.add_argument('--dropout_parity', action='store_true', help="draw dropout masks from torch's CPU generator like the reference instead of in-kernel Philox")
    return parser


def validate(args):
    """The reference's flag exclusivity checks (:169-174) and derived fields (:176)."""
    if args.concat and not args.sisa:
        raise ValueError("The --concat option can only be used with the --sisa option.")
    if args.vanilla and (args.sisa or args.concat):
        raise ValueError("The --vanilla option cannot be used with --sisa or --concat.")
    if args.control and (args.vanilla or args.sisa or args.concat):
        raise ValueError("The --control option cannot be used with any other options.")
    args.client_num_in_total = args.world_size - 1
    return args


def place_entities(args):
    """Bob on cuda:0; Alice r on cuda:(1 + (r-1) mod (n-1)) when the box shows n > 1 GPUs (all on
    cuda:0 otherwise).  `--n_gpus` limits how many of the visible GPUs are used.  Peer access is
    enabled between all of them: the entity classes let a kernel on one GPU store its result in
    another GPU's memory."""
    n = torch.cuda.device_count()
    if n == 0:
        from ._lib import SplitNNError
        raise SplitNNError("no CUDA device: the split-NN hot path has no CPU fallback")
    want = getattr(args, "n_gpus", None)
    n = max(1, min(n, want)) if want else n
    args.bob_device = torch.device("cuda", 0)
    args.n_gpus = n
    args.alice_devices = {r: torch.device("cuda", 0 if n == 1 else 1 + (r - 1) % (n - 1))
                          for r in range(1, args.client_num_in_total + 1)}
    if n > 1:
        from . import _lib
        _lib.check(_lib.load().sn_p2p_enable_all(n))
    return args


def example(rank, world_size, args):
    """The body the reference runs on rank 0 (:42-135).  `rank` is kept for signature
    compatibility; ranks > 0 have nothing to do in a single-process run."""
    if rank != 0:
        return None
    _, bob = get_alice_and_bob(args)
    BOB = bob(args)
    unlearn_request_from_alices = [1]
    omit_label = 9

    if not args.control:
        if not args.sisa:
            for it in range(args.iterations):
                for client_id in range(1, world_size):
                    print(f"Training client {client_id}")
                    BOB.train_request(client_id)
                BOB.eval_request_breakdown(omit_label)
            # Test on the control group
            for it in range(args.iterations):
                BOB.unlearn_request(client_id=unlearn_request_from_alices[0], omit_label=omit_label)
                BOB.eval_request_breakdown(omit_label)
        else:
            print("Training all clients in parallel")
            BOB.train_request_parallel()
            BOB.freeze_alice_weights(range(1, args.client_num_in_total + 1))
            print("Training server")
            BOB.train_and_backward(unlearn_request_from_alices=[], unlearn_id=None)
            BOB.eval_request_breakdown(omit_label)
            # The reference's --concat branch stops here (:64-72); --sisa continues with unlearning
            # (:95-110).  This build runs the unlearning phase for both (BASELINE.json config 4).
            BOB.unfreeze_alice_weights(unlearn_request_from_alices)
            print(f"Retraining client {unlearn_request_from_alices}")
            BOB.unlearn_request(client_id=unlearn_request_from_alices[0], omit_label=omit_label)
            BOB.freeze_alice_weights(unlearn_request_from_alices)
            print("Retraining server upon the omitted labels")
            BOB.train_and_backward(unlearn_request_from_alices, unlearn_id=omit_label)
            BOB.eval_request_breakdown(omit_label)
    else:
        # The reference selects data_entities.py for --control (:20-21), whose bob has no
        # train_request_control (AttributeError at :126).  The control-group methods live in the
        # SISA module, so that is the module used here.
        from .data_entities_vanilla_sisa import bob as sisa_bob
        BOB = sisa_bob(args)
        for it in range(args.iterations):
            for client_id in range(1, world_size):
                print(f"(Control group) Training client {client_id}")
                if client_id in unlearn_request_from_alices:
                    BOB.train_request_control(client_id, omit_label)
                else:
                    BOB.train_request(client_id)
        BOB.freeze_alice_weights(range(1, args.client_num_in_total + 1))
        for it in range(args.iterations):
            print("(Control group) Training server")
            BOB.train_and_backward(unlearn_request_from_alices, unlearn_id=omit_label)
            BOB.eval_request()
    torch.cuda.synchronize()
    return BOB


def main(argv=None):
    args = validate(build_parser().parse_args(argv))
    if args.seed is not None:
        import random
        import numpy as np
        np.random.seed(args.seed)
        random.seed(args.seed)
        args.torch_seeds = {"bob": 100, "alice_base": 200}
    os.makedirs("logs", exist_ok=True)
    place_entities(args)
    from .data.mnist_flat.mnist_flat_generator import load_mnist_image
    load_mnist_image(args)                      # every invocation, in the parent, as the reference (:179)
    return example(0, args.world_size, args)


if __name__ == "__main__":
    main()
