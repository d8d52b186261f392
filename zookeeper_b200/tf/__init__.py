from .dataset import ArrayDataset, Dataset, SyntheticDataset, TFDSDataset
from .experiment import Experiment
from .preprocessing import Preprocessing

__all__ = ["ArrayDataset", "Dataset", "Experiment", "Preprocessing", "SyntheticDataset", "TFDSDataset"]
