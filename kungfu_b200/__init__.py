"""kungfu_b200: adaptive data-parallel training for 8xB200 (NVLink 5 / NVSwitch).

A from-scratch, Blackwell-native library with the capabilities and public API
of KungFu (reference: /root/reference/README.md): the ``kungfu-run`` launcher,
``kungfu_b200.torch`` optimizer wrappers (S-SGD, SMA, PairAveraging, monitors),
broadcast / barrier / consensus / resize_cluster, adaptive strategy selection.
"""
__version__ = '0.1.0'
