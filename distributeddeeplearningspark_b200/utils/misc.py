"""Small helpers mirroring ``distkeras.utils`` pieces the scripts touch
(``get_os_username`` at /root/reference/ddl_mnist_aztk.py:71) plus env knobs."""
from __future__ import annotations

import getpass
import os
import socket


def get_os_username() -> str:
    try:
        return getpass.getuser()
    except Exception:  # no passwd entry in minimal containers
        return os.environ.get('USER', 'unknown')


def determine_host_address() -> str:
    """Address other ranks can reach this host at; loopback when the hostname does not resolve."""
    try:
        return socket.gethostbyname(socket.gethostname())
    except OSError:
        return '127.0.0.1'


def free_port(host: str = '127.0.0.1') -> int:
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind((host, 0))
        return s.getsockname()[1]


def env_int(name: str, default: int) -> int:
    v = os.environ.get(name)
    return default if v in (None, '') else int(v)


def env_str(name: str, default: str) -> str:
    v = os.environ.get(name)
    return default if v in (None, '') else v


def env_flag(name: str, default: bool = False) -> bool:
    v = os.environ.get(name)
    if v in (None, ''):
        return default
    return v.lower() not in ('0', 'false', 'no', 'off')
