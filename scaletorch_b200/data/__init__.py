from .dataloader import MicroBatchDataLoader, SyntheticTokens
from .dataset import DatasetProcessor, get_tokenize_strategy, register_tokenize_strategy
from .pretrain_dataset import PretrainDataset

__all__ = ["DatasetProcessor", "MicroBatchDataLoader", "PretrainDataset", "SyntheticTokens", "get_tokenize_strategy",
           "register_tokenize_strategy"]
