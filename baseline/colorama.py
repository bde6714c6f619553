"""Minimal stand-in for the ``colorama`` package (not installed in this image, not in the offline wheelhouse).

The reference imports ``Fore`` / ``Style`` in ``scaletorch/utils/logger_utils.py:13`` only to colour log lines.
This module is on ``sys.path`` solely for the ``--impl reference`` benchmark arm; it changes nothing in the
reference's code path.
"""


class _Codes:
    def __getattr__(self, name: str) -> str:
        return ""


Fore = _Codes()
Back = _Codes()
Style = _Codes()


def init(*args, **kwargs) -> None:
    return None
