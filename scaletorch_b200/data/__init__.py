from .dataloader import MicroBatchDataLoader, SyntheticTokens
from .dataset import DatasetProcessor, get_tokenize_strategy, register_tokenize_strategy
from .native_loader import NativeTokenLoader, native_loader_available
from .pretrain_dataset import PretrainDataset

__all__ = ["DatasetProcessor", "MicroBatchDataLoader", "NativeTokenLoader", "native_loader_available", "PretrainDataset", "SyntheticTokens", "get_tokenize_strategy",
           "register_tokenize_strategy"]
