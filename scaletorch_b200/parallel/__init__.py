from .context import ParallelContext
from .process_group import (ProcessGroupManager, get_process_group_manager, process_group_manager,
                            setup_process_group_manager)

__all__ = ["ParallelContext", "ProcessGroupManager", "get_process_group_manager", "process_group_manager",
           "setup_process_group_manager"]
