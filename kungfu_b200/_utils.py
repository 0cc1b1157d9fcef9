"""Small helpers with the names and return conventions of the reference's private
Python module (/root/reference/srcs/python/kungfu/_utils.py): note that ``measure``
here returns ``(seconds, result)``, the reverse of ``kungfu_b200.utils.measure``
(which follows the Go runtime's ``utils.Measure``)."""
import time

from .utils import (log_event, map_maybe, one_based_range, show_duration,  # noqa: F401
                    since_job_start, since_proc_start)


def measure(f):
    t0 = time.time()
    result = f()
    return time.time() - t0, result


_since_job_start, _since_proc_start, _log_event = since_job_start, since_proc_start, log_event
