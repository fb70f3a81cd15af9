"""Storage providers (reference: LinqToDryad/DataProvider.cs, DataPath.cs)."""
import pytest

from dryad_b200.api.context import DryadLinqContext


def test_data_provider_registration(tmp_path):
    """A new storage scheme is one DataProvider subclass + register (DataProvider.cs:166-183)."""
    from dryad_b200.channel import store
    from dryad_b200.channel.provider import DataProvider, MemoryDataProvider
    from dryad_b200.api.errors import DryadLinqException
    assert {"partfile", "file"} <= set(DataProvider.schemes())
    with pytest.raises(DryadLinqException):
        store.exists("mem://t1")                          # nobody serves mem:// yet
    with pytest.raises(DryadLinqException):
        store.exists("hdfs://namenode/x")                 # cluster file systems need a registered provider
    mem = MemoryDataProvider()
    DataProvider.register(mem)
    try:
        with pytest.raises(DryadLinqException):
            DataProvider.register(MemoryDataProvider())   # a scheme is registered once
        ctx = DryadLinqContext(3, work_dir=str(tmp_path))
        q = ctx.from_enumerable(range(30), 3).select(lambda x: x * 2).to_store("mem://doubled")
        ctx.submit_and_wait(q)
        assert store.exists("mem://doubled")
        assert store.stream_info("mem://doubled").partition_count == 3
        back = sorted(ctx.from_store("mem://doubled"))
        assert back == [x * 2 for x in range(30)]
        with pytest.raises(DryadLinqException):           # existing stream, no delete_if_exists
            ctx.submit_and_wait(ctx.from_enumerable([1]).to_store("mem://doubled"))
        store.delete("mem://doubled")
        assert not store.exists("mem://doubled")
    finally:
        DataProvider.unregister("mem")
