"""Console entry points (pip would install `kungfu-run` -> kungfu_b200.cmd:run).
Reference: /root/reference/srcs/python/kungfu/cmd/__init__.py:4-6 -> GoKungfuRunMain."""
import ctypes as C
import sys

from .. import _host


def _main(symbol, argv):
    l = _host.lib()
    fn = getattr(l, symbol)
    fn.argtypes = [C.c_int, C.POINTER(C.c_char_p)]
    fn.restype = C.c_int
    args = (C.c_char_p * len(argv))(*[a.encode() for a in argv])
    return fn(len(argv), args)


def run():
    sys.exit(_main('kf_run_main', ['kungfu-run'] + sys.argv[1:]))


def config_server():
    sys.exit(_main('kf_config_server_main', ['kungfu-config-server'] + sys.argv[1:]))
