"""B200-native distributed k-NN classification engine (Python layer).

Sub-packages: ``ops`` (CPU oracle + sm_100a kernels), ``models`` (the k-NN
classifier, the framework's flagship "model"), ``parallel`` (process grid,
symmetric peer memory, fused and NCCL baseline engines), ``utils`` (input
format, timing, clocks).
"""
__version__ = "0.1.0"
