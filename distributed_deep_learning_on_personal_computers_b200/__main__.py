"""``python -m distributed_deep_learning_on_personal_computers_b200 [--flags]`` -- run a training job.

One process per GPU (or per CPU rank).  Launch N ranks on one box with::

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29500 -m distributed_deep_learning_on_personal_computers_b200 --epochs 2

or start ranks by hand with ``--rank R --world-size N --master-addr A --master-port P`` (the
reference's model: one script per PC, rank 0 is the server, ``R:690`` / ``R:792``).
"""
from .launch import main

if __name__ == "__main__":
    main()
