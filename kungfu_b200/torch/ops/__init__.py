from .collective import (all_gather, all_reduce_fn, barrier, broadcast,  # noqa: F401
                         broadcast_parameters, consensus, inplace_all_reduce_async_op,
                         inplace_all_reduce_op, inplace_broadcast_async_op,
                         monitored_all_reduce, request_variable, save_variable,
                         wait_all_handles, wait_handle)
