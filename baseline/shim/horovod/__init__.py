"""Stand-in for Horovod (not installable offline: no wheel, needs MPI/NCCL toolchains and an index).

Only the ``horovod.torch`` symbols the reference touches are provided, mapped 1:1 onto
``torch.distributed`` collectives (NCCL on GPU, gloo on CPU), so the UNMODIFIED reference code runs its own
algorithm - all-gather + ``torch.matmul`` loops, N all-reduces for ``tn`` - on this box."""
__version__ = '0.28.1'
