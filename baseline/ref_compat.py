"""Environment shims so the UNMODIFIED reference (DeepSpeed 0.12.7, installed under
baseline/_ref) imports on torch 2.11 / Python 3.12.  Nothing in baseline/_ref is edited;
we only re-create names that newer torch removed, *before* `import deepspeed`.
"""
import logging
import os
import sys


def install_shims():
    os.environ.setdefault("DS_SKIP_CUDA_CHECK", "1")
    here = os.path.dirname(os.path.abspath(__file__))
    ref = os.path.join(here, "_ref")
    if ref not in sys.path:
        sys.path.insert(0, ref)
    import torch  # noqa
    # torch.distributed.elastic.agent.server.api lost `log` / `_get_socket_with_port`
    try:
        import torch.distributed.elastic.agent.server.api as api
        if not hasattr(api, "log"):
            api.log = logging.getLogger("torch.distributed.elastic.agent.server.api")
        if not hasattr(api, "_get_socket_with_port"):
            import socket

            def _get_socket_with_port():
                s = socket.socket(socket.AF_INET, socket.SOCK_STREAM)
                s.bind(("127.0.0.1", 0))
                s.listen(0)
                return s

            api._get_socket_with_port = _get_socket_with_port
    except Exception:  # pragma: no cover
        pass
    # numpy 2 dropped the BUFSIZE constant (imported, unused, by autotuning/scheduler.py)
    import numpy
    if not hasattr(numpy, "BUFSIZE"):
        numpy.BUFSIZE = 8192
    return ref
