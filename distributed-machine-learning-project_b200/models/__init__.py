"""Model family: brute-force k-NN classification (the reference's only workload)."""
from .knn import KNNClassifier, KNNResult, choose_path, knn  # noqa: F401
