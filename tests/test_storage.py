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


def test_user_types_are_serialized_automatically(tmp_path):
    """Records of user classes need no hand-written serializer: public fields are written by name (dataclasses,
    plain classes, __slots__, inheritance, nesting, lists of objects, enums); circular references and classes
    defined inside a function are errors with their own codes (SerializationTests.cs, DryadLinqCodeGen.cs:496-1479)."""
    import serialization_types as T
    from dryad_b200.api.errors import DryadLinqException, ErrorCode
    from dryad_b200.channel import records as R
    objs = [T.Point(1, 2.5), T.Frozen("f", ("a", "b")), T.Nested(T.Point(3, 4.0), [T.Point(5, 6.0), T.Point(7, 8.0)], T.Color.GREEN, T.Plain(1, [2])),
            T.Plain(9, [T.Slotted(1, "x")]), T.Slotted((1, 2), None), T.Derived(1, [2], "z"), T.Color.RED, [T.Color.RED, T.Point(0, 0.0)]]
    for o in objs:
        back, pos = R.decode(memoryview(R.encode_record(o)), 0)
        assert back == o and type(back) is type(o), o
    assert R.decode_many(R.encode_many(objs)[0]) == objs                    # through the native codec's hand-back path
    a = T.Node(1)
    a.next = T.Node(2)
    assert R.decode(memoryview(R.encode_record(a)), 0)[0].next.v == 2       # a chain is fine
    a.next.next = a
    with pytest.raises(DryadLinqException) as e:
        R.encode_record(a)
    assert e.value.error_code == ErrorCode.CannotHandleCircularTypes

    class Local:
        def __init__(self):
            self.q = 1
    with pytest.raises(DryadLinqException) as e:
        R.encode_record(Local())
    assert e.value.error_code == ErrorCode.TypeRequiredToBePublic
    with pytest.raises(DryadLinqException) as e:
        R.encode_record(lambda x: x)
    assert e.value.error_code == ErrorCode.TypeNotSerializable
    # end to end: a table of user objects written and read back, and carried through a shuffle
    ctx = DryadLinqContext(3, work_dir=str(tmp_path))
    uri = f"partfile://{tmp_path}/points.pt"
    pts = [T.Point(i, i * 0.5) for i in range(50)]
    ctx.submit_and_wait(ctx.from_enumerable(pts, 3).hash_partition(lambda p: p.x, 2).to_store(uri))
    assert sorted(ctx.from_store(uri), key=lambda p: p.x) == pts
