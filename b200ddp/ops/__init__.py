"""Python faces of the sm_100a kernels (see ``csrc/`` and ``include/b200ddp.h``)."""
