"""dryad_b200 — a B200-native dataflow engine with the capabilities of Dryad / DryadLINQ.

Public API (see dryad_b200.api): DryadLinqContext, Query operators, Decomposable / Associative
protocols, record types; column expressions in dryad_b200.expr.
"""
from .api.context import (CompressionScheme, DryadLinqB200Cluster, DryadLinqCluster, DryadLinqContext, DryadLinqJobInfo,
                          DryadLinqLocalCluster, JobStatus, PlatformKind, QueryLoggingLevel, Table)
from .channel.provider import DataProvider
from .api.decomposable import IAssociative, IDecomposable, associative, decomposable, resource
from .api.errors import DryadLinqException, ErrorCategory, ErrorCode
from .api.query import KeyedMultiQuery, MultiQuery, Query
from .api.rows import RowKey
from .api.types import AggregateValue, ForkTuple, ForkValue, Grouping, LineRecord, Pair, Record
from .expr import col, lit, where

__version__ = "0.2.1b200"
__all__ = [
    "DryadLinqContext", "DryadLinqJobInfo", "JobStatus", "PlatformKind", "QueryLoggingLevel", "CompressionScheme",
    "DryadLinqCluster", "DryadLinqLocalCluster", "DryadLinqB200Cluster", "DataProvider", "Table", "IDecomposable", "IAssociative", "decomposable", "associative", "resource", "DryadLinqException",
    "ErrorCode", "ErrorCategory", "Query", "MultiQuery", "KeyedMultiQuery", "Record", "Pair", "LineRecord",
    "RowKey", "ForkTuple", "ForkValue", "Grouping", "AggregateValue", "col", "lit", "where",
]
