"""dmlc_core_b200: B200-native text -> CSR RowBlock parsers (LibSVM / LibFM / CSV).

Drop-in for the hot path of rahul003/dmlc-core (`dmlc::Parser<IndexType>`); see DESIGN.md.
The compute path is hand-written sm_100a CUDA behind the C ABI of include/dmlc_b200_parser.h.
"""
from . import capi  # noqa: F401
from .capi import Parser, ParseError  # noqa: F401

__all__ = ["capi", "Parser", "ParseError"]
