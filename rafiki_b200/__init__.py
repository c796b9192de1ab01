"""rafiki_b200 -- B200-native serving hot path of Rafiki (predict -> ensemble).

Sub-packages mirror the reference's layout for the path they replace:
``model`` (plugin API), ``predictor``, ``worker``, ``cache``; ``engine`` is the host runtime
over ``librafiki_b200.so`` (``csrc/``, declared in ``include/rafiki_b200.h``).
"""
__version__ = '0.1.0'
