"""speech_anonymization_b200 -- B200-native (sm_100a) Fbank + anonymizer hot path of
viswavi/speech-anonymization behind the reference's own Python class API.

Everything numeric runs in libspanon.so (include/spanon.h); there is no CPU or PyTorch
fallback -- a missing library or a non-CUDA tensor raises.
"""
from ._lib import SpanonError, lib  # noqa: F401
from .features import Fbank, InputNormalization, pad_time  # noqa: F401
from .losses import (ClusterMI, CosineSimilarityLoss, GroupSampler, GroupSamplingMI, MILoss,  # noqa: F401
                     SimilarityMetricsStats, cosine_distance_2d, knn_mi)
from .models import (GLU, ConvAutoencoder, ConvReconstruction, FullyConnectedAutoencoder, FullyConnSexClassifier,  # noqa: F401
                     GradReverse, StatisticsPooling, TDNNSexClassifier, Xvector, XvectorClassifier, XvectorHead)
from .augment import SpecAugment  # noqa: F401
from .distributed import DataParallelTrainer, FlatGradients, gather_for_mi  # noqa: F401
from .pipeline import AnonymizerPipeline, GraphedStep, StreamPipe, shard_range  # noqa: F401

__all__ = ["Fbank", "InputNormalization", "ConvAutoencoder", "ConvReconstruction", "XvectorClassifier", "FullyConnectedAutoencoder", "CosineSimilarityLoss",
           "MILoss", "ClusterMI", "GroupSamplingMI", "AnonymizerPipeline", "StreamPipe", "shard_range", "pad_time", "lib"]
