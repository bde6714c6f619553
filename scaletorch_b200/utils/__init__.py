from .device import (get_current_device, get_device_type, get_dist_backend, get_dist_info, get_theoretical_flops,
                     is_bf16_supported, register_device_flops, set_device)
from .env_utils import cleanup_dist, get_system_info, init_dist_pytorch
from .logger_utils import get_logger
from .misc import (assert_no_meta_tensors, average_loss_across_dp_cp_ranks, get_mfu, get_num_params,
                   rank_print, set_all_seed, to_readable_format)
from .monitor import PerformanceMonitor
from .path import check_file_exist, mkdir_or_exist

__all__ = [
    "PerformanceMonitor", "assert_no_meta_tensors", "average_loss_across_dp_cp_ranks", "check_file_exist",
    "cleanup_dist", "get_dist_info", "get_system_info", "init_dist_pytorch", "get_current_device", "get_device_type", "get_dist_backend", "get_logger", "get_mfu", "get_num_params",
    "get_theoretical_flops", "is_bf16_supported", "mkdir_or_exist", "rank_print", "register_device_flops",
    "set_all_seed", "set_device", "to_readable_format",
]
