"""Small path helpers (reference: ``scaletorch/utils/path.py``)."""
import os


def check_file_exist(filename: str, msg_tmpl: str = 'file "{}" does not exist') -> None:
    """Raise ``FileNotFoundError`` (message from ``msg_tmpl``) unless ``filename`` is an existing file."""
    if not os.path.isfile(filename):
        raise FileNotFoundError(msg_tmpl.format(filename))


def mkdir_or_exist(dir_name: str, mode: int = 0o777) -> None:
    """``os.makedirs(dir_name, exist_ok=True)`` with ``~`` expanded; an empty name is ignored."""
    if dir_name == "":
        return
    os.makedirs(os.path.expanduser(dir_name), mode=mode, exist_ok=True)
