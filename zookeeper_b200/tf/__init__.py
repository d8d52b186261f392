from .dataset import ArrayDataset, Dataset, MultiTFDSDataset, SyntheticDataset, TFDSDataset
from .experiment import Experiment
from .preprocessing import Preprocessing

__all__ = ["ArrayDataset", "Dataset", "Experiment", "MultiTFDSDataset", "Preprocessing", "SyntheticDataset",
           "TFDSDataset"]
