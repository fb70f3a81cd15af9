"""Typed error surface (reference: LinqToDryad/DryadLinqException.cs:34,
LinqToDryad/DryadLinqFaultCodes.cs:31-278).  Codes keep the reference's category layout
(category << 24 | ordinal) and the same names, so callers can switch on them."""
from __future__ import annotations

from enum import IntEnum


class ErrorCategory(IntEnum):
    QueryAPI = 0x01000000
    CodeGen = 0x02000000
    JobSubmission = 0x03000000
    Serialization = 0x04000000
    StoreClient = 0x05000000
    VertexRuntime = 0x06000000
    LocalDebug = 0x07000000
    Unknown = 0x0F000000


_NAMES = {
    ErrorCategory.CodeGen: [
        "TypeRequiredToBePublic", "CustomSerializerMustSupportDefaultCtor", "CustomSerializerMustBeClassOrStruct",
        "TypeNotSerializable", "CannotHandleSubtypes", "UDTMustBeConcreteType", "UDTHasFieldOfNonPublicType",
        "UDTIsDelegateType", "FailedToBuild", "OutputTypeCannotBeAnonymous", "InputTypeCannotBeAnonymous",
        "BranchOfForkNotUsed", "ComparerMustBeSpecifiedOrKeyTypeMustBeIComparable",
        "ComparerMustBeSpecifiedOrKeyTypeMustBeIEquatable",
        "ComparerExpressionMustBeSpecifiedOrElementTypeMustBeIEquatable", "TooManyHomomorphicAttributes",
        "HomomorphicApplyNeedsSamePartitionCount", "UnrecognizedDataSource",
        "CannotConcatDatasetsWithDifferentCompression", "AggregateOperatorNotSupported",
        "FinalizerReturnTypeMismatch", "CannotHandleCircularTypes", "OperatorNotSupported",
        "AggregationOperatorRequiresIComparable", "DecomposerTypeDoesNotImplementInterface",
        "DecomposerTypeImplementsTooManyInterfaces", "DecomposerTypesDoNotMatch", "DecomposerTypeMustBePublic",
        "DecomposerTypeDoesNotHavePublicDefaultCtor", "AssociativeMethodHasWrongForm",
        "AssociativeTypeDoesNotImplementInterface", "AssociativeTypeImplementsTooManyInterfaces",
        "AssociativeTypesDoNotMatch", "AssociativeTypeMustBePublic", "AssociativeTypeDoesNotHavePublicDefaultCtor",
        "CannotCreatePartitionNodeRandom", "PartitionKeysNotProvided", "PartitionKeysAreNotConsistentlyOrdered",
        "IsDescendingIsInconsistent", "BadSeparatorCount", "TypeMustHaveDataMembers", "CannotHandleObjectFields",
        "CannotHandleDerivedtypes", "MultipleOutputsWithSameDscUri", "OutputUriAlsoQueryInput", "Internal",
    ],
    ErrorCategory.StoreClient: [
        "DSCStreamError", "StreamDoesNotExist", "StreamAlreadyExists", "AttemptToReadFromAWriteStream",
        "FailedToCreateStream", "JobToCreateTableWasCanceled", "FailedToGetReadPathsForStream",
        "CannotAccesFilePath", "PositionNotSupported", "GetFileSizeError", "ReadFileError",
        "UnknownCompressionScheme", "WriteFileError", "MultiBlockEmptyPartitionList", "GetURINotSupported",
        "SetCalcFPNotSupported", "GetFPNotSupported", "FailedToAllocateNewNativeBuffer",
        "FailedToReadFromInputChannel", "FailedToWriteToOutputChannel", "MultiBlockCannotAccesFilePath",
    ],
    ErrorCategory.JobSubmission: [
        "DryadHomeMustBeSpecified", "ClusterNameMustBeSpecified", "UnexpectedJobStatus", "JobStatusQueryError",
        "JobOptionNotImplemented", "DryadLinqJobMinMustBe", "SubmissionFailure", "UnsupportedSchedulerType",
        "UnsupportedExecutionKind", "DidNotCompleteSuccessfully", "Binaries",
    ],
    ErrorCategory.QueryAPI: [
        "DistinctAttributeComparerNotDefined", "SerializerTypeMustBeNonNull",
        "SerializerTypeMustSupportIDryadLinqSerializer", "UnrecognizedOperatorName", "UnsupportedExpressionsType",
        "UnsupportedExpressionType", "IndexTooSmall", "MultiQueryableKeyOutOfRange", "IndexOutOfRange",
        "ExpressionTypeNotHandled", "FailedToGetStreamProps", "MetadataRecordType", "JobToCreateTableFailed",
        "OnlyAvailableForPhysicalData", "FileSetMustBeSealed", "FileSetCouldNotBeOpened",
        "FileSetMustHaveAtLeastOneFile", "CouldNotGetClientVersion", "CouldNotGetServerVersion", "ContextDisposed",
        "UnhandledQuery", "ExpressionMustBeMethodCall", "UntypedProviderMethodsNotSupported", "ErrorReadingMetadata",
        "MustStartFromContext",
    ],
    ErrorCategory.Serialization: [
        "FailedToReadFrom", "EndOfStreamEncountered", "SettingPositionNotSupported", "FingerprintDisabled",
        "RecordSizeMax", "ReadByteNotAllowed", "ReadNotAllowed", "SeekNotSupported", "SetLengthNotSupported",
        "FailedToDeserialize", "ChannelCannotBeReadMoreThanOnce", "WriteNotSupported", "WriteByteNotSupported",
        "CannotSerializeDryadLinqQuery", "CannotSerializeObject", "GeneralSerializeFailure",
    ],
    ErrorCategory.VertexRuntime: [
        "SourceOfMergesortMustBeMultiEnumerable", "ThenByNotSupported", "AggregateNoElements", "FirstNoElementsFirst",
        "SingleMoreThanOneElement", "SingleNoElements", "LastNoElements", "MinNoElements", "MaxNoElements",
        "AverageNoElements", "RangePartitionKeysMissing", "PartitionFuncReturnValueExceedsNumPorts",
        "FailureInExcept", "FailureInIntersect", "FailureInSort", "RangePartitionInputOutputMismatch", "KeyNotFound",
        "TooManyItems", "FailureInHashGroupBy", "FailureInSortGroupBy", "FailureInHashJoin",
        "FailureInHashGroupJoin", "FailureInDistinct", "FailureInOperator", "FailureInUserApplyFunction",
        "FailureInOrderedGroupBy", "TooManyElementsBeforeReduction",
    ],
    ErrorCategory.LocalDebug: ["CreatingDscDataFromLocalDebugFailed"],
    ErrorCategory.Unknown: ["UnknownError"],
}


class _Codes:
    def __init__(self):
        self._by_value = {}
        for cat, names in _NAMES.items():
            for i, n in enumerate(names):
                v = int(cat) + i
                setattr(self, n, v)
                self._by_value[v] = n

    def name(self, code: int) -> str:
        return self._by_value.get(code, f"0x{code:08x}")

    def category(self, code: int) -> ErrorCategory:
        return ErrorCategory(code & 0x0F000000)


ErrorCode = _Codes()


class DryadLinqException(Exception):
    """All framework failures surface as this type with a numeric `error_code`."""

    def __init__(self, code: int, message: str = "", inner: BaseException | None = None):
        super().__init__(f"[{ErrorCode.name(code)}] {message}")
        self.error_code = code
        self.message = message
        self.inner = inner

    @property
    def category(self) -> ErrorCategory:
        return ErrorCode.category(self.error_code)
