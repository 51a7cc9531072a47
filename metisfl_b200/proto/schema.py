"""Programmatic protobuf schema for the MetisFL wire contract.

There is no ``protoc`` / ``grpc_tools`` in the build image, so the five schema
files of the reference (``metisfl/proto/{model,metis,controller,learner,
service_common}.proto``) are rebuilt here as ``FileDescriptorProto`` objects
from a small table-driven description, registered in a private descriptor
pool, and turned into concrete message classes with ``message_factory``.

Field names and numbers follow the reference contract exactly (see SURVEY.md
§9.1; reference ``metisfl/proto/model.proto:14-152``, ``metis.proto:13-365``,
``controller.proto:8-176``, ``learner.proto:8-64``,
``service_common.proto:7-31``) so that bytes produced by either side parse on
the other.  Extensions use fresh field numbers only:

* ``DType.Type``: ``FLOAT16 = 10``, ``BFLOAT16 = 11``
* ``TensorSpec.device_slot = 15`` (``DeviceSlotRef``): a control-only reference
  to a tensor that already lives in a peer-visible HBM slot.
* ``AggregationRule.fed_async = 6`` (``FedAsync``): staleness-weighted
  asynchronous mixing rule.
* ``ProtocolSpecs.task_deadline_secs = 15``: controller-side eviction of
  learners that never report (fixes reference quirk Q9).
* ``JoinFederationRequest.fabric = 15`` (``FabricMembership``): a learner that
  hosts a rank of the on-box NVLink fabric says so when it joins.
* ``LearnerService.AggregateOnDevice``: the controller's command to the GPU
  processes to run the aggregation kernels over the device slots it names
  (the reference has no counterpart: its controller sums protobuf bytes on the
  CPU, ``controller.cc:795-950``).
"""
from __future__ import annotations

from google.protobuf import descriptor_pb2 as dpb
from google.protobuf import descriptor_pool, message_factory
from google.protobuf import timestamp_pb2  # noqa: F401  (registers Timestamp)

PKG = "metisfl"
_FD = dpb.FieldDescriptorProto

_SCALARS = {
    "double": _FD.TYPE_DOUBLE, "float": _FD.TYPE_FLOAT, "int64": _FD.TYPE_INT64,
    "uint64": _FD.TYPE_UINT64, "int32": _FD.TYPE_INT32, "uint32": _FD.TYPE_UINT32,
    "bool": _FD.TYPE_BOOL, "string": _FD.TYPE_STRING, "bytes": _FD.TYPE_BYTES,
}
TS = ".google.protobuf.Timestamp"


# --------------------------------------------------------------------------
# tiny schema DSL
# --------------------------------------------------------------------------
class _Field:
    def __init__(self, name, number, typ, repeated=False, oneof=None,
                 optional=False, map_of=None):
        self.name, self.number, self.typ = name, number, typ
        self.repeated, self.oneof, self.optional, self.map_of = repeated, oneof, optional, map_of


def f(name, number, typ, **kw):
    """Singular field. ``typ`` is a scalar name, ``"enum:Path"`` or a message path."""
    return _Field(name, number, typ, **kw)


def rep(name, number, typ):
    return _Field(name, number, typ, repeated=True)


def mp(name, number, ktyp, vtyp):
    """``map<ktyp, vtyp>`` field."""
    return _Field(name, number, None, repeated=True, map_of=(ktyp, vtyp))


class _Enum:
    def __init__(self, name, **values):
        self.name, self.values = name, values


class _Msg:
    def __init__(self, name, *items):
        self.name, self.items = name, items


def _camel(s):
    return "".join(p[:1].upper() + p[1:] for p in s.split("_"))


def _set_type(fd, typ):
    if typ in _SCALARS:
        fd.type = _SCALARS[typ]
    elif typ.startswith("enum:"):
        fd.type = _FD.TYPE_ENUM
        fd.type_name = _qual(typ[5:])
    else:
        fd.type = _FD.TYPE_MESSAGE
        fd.type_name = _qual(typ)


def _qual(path):
    return path if path.startswith(".") else f".{PKG}.{path}"


def _build_msg(spec: _Msg, scope: str) -> dpb.DescriptorProto:
    d = dpb.DescriptorProto(name=spec.name)
    full = f"{scope}.{spec.name}"
    oneofs = {}
    synthetic = []
    for it in spec.items:
        if isinstance(it, _Enum):
            e = d.enum_type.add(name=it.name)
            for k, v in it.values.items():
                e.value.add(name=k, number=v)
        elif isinstance(it, _Msg):
            d.nested_type.append(_build_msg(it, full))
        elif isinstance(it, _Field):
            fd = d.field.add(name=it.name, number=it.number)
            fd.label = _FD.LABEL_REPEATED if it.repeated else _FD.LABEL_OPTIONAL
            if it.map_of is not None:
                entry = d.nested_type.add(name=_camel(it.name) + "Entry")
                entry.options.map_entry = True
                kf = entry.field.add(name="key", number=1, label=_FD.LABEL_OPTIONAL)
                _set_type(kf, it.map_of[0])
                vf = entry.field.add(name="value", number=2, label=_FD.LABEL_OPTIONAL)
                _set_type(vf, it.map_of[1])
                fd.type = _FD.TYPE_MESSAGE
                fd.type_name = f".{full.lstrip('.')}.{entry.name}"
            else:
                _set_type(fd, it.typ)
            if it.oneof is not None:
                if it.oneof not in oneofs:
                    oneofs[it.oneof] = len(d.oneof_decl)
                    d.oneof_decl.add(name=it.oneof)
                fd.oneof_index = oneofs[it.oneof]
            elif it.optional:
                synthetic.append(fd)
    # proto3 `optional` = synthetic oneofs, which must come after the real ones.
    for fd in synthetic:
        fd.proto3_optional = True
        fd.oneof_index = len(d.oneof_decl)
        d.oneof_decl.add(name="_" + fd.name)
    return d


def _build_file(name, deps, msgs, services=()):
    fdp = dpb.FileDescriptorProto(name=name, package=PKG, syntax="proto3")
    fdp.dependency.extend(deps)
    for m in msgs:
        fdp.message_type.append(_build_msg(m, "." + PKG))
    for sname, methods in services:
        s = fdp.service.add(name=sname)
        for mname in methods:
            s.method.add(name=mname, input_type=_qual(mname + "Request"),
                         output_type=_qual(mname + "Response"))
    return fdp


# --------------------------------------------------------------------------
# model.proto
# --------------------------------------------------------------------------
MODEL = "metisfl/proto/model.proto"
_model_msgs = [
    _Msg("DType",
         _Enum("Type", INT8=0, INT16=1, INT32=2, INT64=3, UINT8=4, UINT16=5, UINT32=6,
               UINT64=7, FLOAT32=8, FLOAT64=9, FLOAT16=10, BFLOAT16=11),
         _Enum("ByteOrder", NA=0, BIG_ENDIAN_ORDER=1, LITTLE_ENDIAN_ORDER=2),
         f("type", 1, "enum:DType.Type"),
         f("byte_order", 2, "enum:DType.ByteOrder"),
         f("fortran_order", 3, "bool")),
    _Msg("TensorQuantifier",
         f("tensor_non_zeros", 1, "uint32", optional=True),
         f("tensor_zeros", 2, "uint32", optional=True),
         f("tensor_size_bytes", 3, "uint32")),
    _Msg("DeviceSlotRef",  # extension: tensor lives in a peer-visible HBM slot
         f("device_rank", 1, "uint32"),
         f("slot", 2, "uint32"),
         f("offset_bytes", 3, "uint64"),
         f("version", 4, "uint64")),
    _Msg("TensorSpec",
         f("length", 1, "uint32"),
         rep("dimensions", 2, "int64"),
         f("type", 3, "DType"),
         f("value", 4, "bytes"),
         f("device_slot", 15, "DeviceSlotRef")),
    _Msg("PlaintextTensor", f("tensor_spec", 1, "TensorSpec")),
    _Msg("CiphertextTensor", f("tensor_spec", 1, "TensorSpec")),
    _Msg("Model",
         _Msg("Variable",
              f("name", 1, "string"),
              f("trainable", 2, "bool"),
              f("plaintext_tensor", 3, "PlaintextTensor", oneof="tensor"),
              f("ciphertext_tensor", 4, "CiphertextTensor", oneof="tensor")),
         rep("variables", 1, "Model.Variable")),
    _Msg("FederatedModel",
         f("num_contributors", 1, "uint32"),
         f("global_iteration", 2, "uint32"),
         f("model", 3, "Model")),
    _Msg("OptimizerConfig",
         f("vanilla_sgd", 1, "VanillaSGD", oneof="config"),
         f("momentum_sgd", 2, "MomentumSGD", oneof="config"),
         f("fed_prox", 3, "FedProx", oneof="config"),
         f("adam", 4, "Adam", oneof="config"),
         f("adam_weight_decay", 5, "AdamWeightDecay", oneof="config")),
    _Msg("VanillaSGD", f("learning_rate", 1, "float"), f("L1_reg", 2, "float"), f("L2_reg", 3, "float")),
    _Msg("MomentumSGD", f("learning_rate", 1, "float"), f("momentum_factor", 2, "float")),
    _Msg("FedProx", f("learning_rate", 1, "float"), f("proximal_term", 2, "float")),
    _Msg("Adam", f("learning_rate", 1, "float"), f("beta_1", 2, "float"),
         f("beta_2", 3, "float"), f("epsilon", 4, "float")),
    _Msg("AdamWeightDecay", f("learning_rate", 1, "float"), f("weight_decay", 2, "float")),
]

# --------------------------------------------------------------------------
# service_common.proto
# --------------------------------------------------------------------------
COMMON = "metisfl/proto/service_common.proto"
_common_msgs = [
    _Msg("Ack", f("status", 1, "bool"), f("timestamp", 2, TS), f("message", 3, "string")),
    _Msg("GetServicesHealthStatusRequest"),
    _Msg("GetServicesHealthStatusResponse", mp("services_status", 1, "string", "bool")),
    _Msg("ShutDownRequest"),
    _Msg("ShutDownResponse", f("ack", 1, "Ack")),
]


def _dataset_spec():
    reg = [f(n, i + 1, "double") for i, n in
           enumerate(["min", "max", "mean", "median", "mode", "stddev"])]
    items = [
        f("num_training_examples", 1, "uint32"),
        f("num_validation_examples", 2, "uint32"),
        f("num_test_examples", 3, "uint32"),
        _Msg("ClassificationDatasetSpec", mp("class_examples_num", 1, "uint32", "uint32")),
        _Msg("RegressionDatasetSpec", *reg),
    ]
    num = 4
    for split in ("training", "validation", "test"):
        items.append(f(f"{split}_classification_spec", num,
                       "DatasetSpec.ClassificationDatasetSpec", oneof=f"{split}_dataset_spec"))
        items.append(f(f"{split}_regression_spec", num + 1,
                       "DatasetSpec.RegressionDatasetSpec", oneof=f"{split}_dataset_spec"))
        num += 2
    return _Msg("DatasetSpec", *items)


# --------------------------------------------------------------------------
# metis.proto
# --------------------------------------------------------------------------
METIS = "metisfl/proto/metis.proto"
_metis_msgs = [
    _Msg("ServerEntity", f("hostname", 1, "string"), f("port", 2, "uint32"),
         f("ssl_config", 3, "SSLConfig")),
    _Msg("SSLConfigFiles", f("public_certificate_file", 1, "string"),
         f("private_key_file", 2, "string")),
    _Msg("SSLConfigStream", f("public_certificate_stream", 1, "bytes"),
         f("private_key_stream", 2, "bytes")),
    _Msg("SSLConfig", f("enable_ssl", 1, "bool"),
         f("ssl_config_files", 6, "SSLConfigFiles", oneof="config"),
         f("ssl_config_stream", 7, "SSLConfigStream", oneof="config")),
    _dataset_spec(),
    _Msg("LearningTaskTemplate", f("num_local_updates", 1, "uint32")),
    _Msg("LearningTask",
         f("global_iteration", 1, "uint32"),
         f("num_local_updates", 2, "uint32"),
         f("training_dataset_percentage_for_stratified_validation", 3, "float"),
         f("metrics", 4, "EvaluationMetrics")),
    _Msg("CompletedLearningTask", f("model", 1, "Model"),
         f("execution_metadata", 2, "TaskExecutionMetadata"), f("aux_metadata", 3, "string")),
    _Msg("TaskExecutionMetadata",
         f("global_iteration", 1, "uint32"),
         f("task_evaluation", 2, "TaskEvaluation"),
         f("completed_epochs", 3, "float"),
         f("completed_batches", 4, "uint32"),
         f("batch_size", 5, "uint32"),
         f("processing_ms_per_epoch", 6, "float"),
         f("processing_ms_per_batch", 7, "float")),
    _Msg("TaskEvaluation",
         rep("training_evaluation", 1, "EpochEvaluation"),
         rep("validation_evaluation", 2, "EpochEvaluation"),
         rep("test_evaluation", 3, "EpochEvaluation")),
    _Msg("EpochEvaluation", f("epoch_id", 1, "uint32"), f("model_evaluation", 2, "ModelEvaluation")),
    _Msg("EvaluationMetrics", rep("metric", 1, "string")),
    _Msg("ModelEvaluation", mp("metric_values", 1, "string", "string")),
    _Msg("ModelEvaluations",
         f("training_evaluation", 1, "ModelEvaluation"),
         f("validation_evaluation", 2, "ModelEvaluation"),
         f("test_evaluation", 3, "ModelEvaluation")),
    _Msg("LocalTasksMetadata", rep("task_metadata", 1, "TaskExecutionMetadata")),
    _Msg("CommunityModelEvaluation", f("global_iteration", 1, "uint32"),
         mp("evaluations", 2, "string", "ModelEvaluations")),
    _Msg("Hyperparameters", f("batch_size", 1, "uint32"), f("optimizer", 2, "OptimizerConfig")),
    _Msg("ControllerParams",
         f("server_entity", 1, "ServerEntity"),
         f("global_model_specs", 2, "GlobalModelSpecs"),
         f("communication_specs", 3, "CommunicationSpecs"),
         f("model_store_config", 4, "ModelStoreConfig"),
         _Msg("ModelHyperparams",
              f("batch_size", 1, "uint32"), f("epochs", 2, "uint32"),
              f("optimizer", 3, "OptimizerConfig"), f("percent_validation", 4, "float")),
         f("model_hyperparams", 5, "ControllerParams.ModelHyperparams")),
    _Msg("ModelStoreConfig",
         f("in_memory_store", 1, "InMemoryStore", oneof="config"),
         f("redis_db_store", 2, "RedisDBStore", oneof="config")),
    _Msg("InMemoryStore", f("model_store_specs", 1, "ModelStoreSpecs")),
    _Msg("RedisDBStore", f("model_store_specs", 1, "ModelStoreSpecs"),
         f("server_entity", 2, "ServerEntity")),
    _Msg("NoEviction"),
    _Msg("LineageLengthEviction", f("lineage_length", 1, "uint32")),
    _Msg("ModelStoreSpecs",
         f("no_eviction", 1, "NoEviction", oneof="eviction_policy"),
         f("lineage_length_eviction", 2, "LineageLengthEviction", oneof="eviction_policy")),
    _Msg("AggregationRule",
         f("fed_avg", 1, "FedAvg", oneof="rule"),
         f("fed_stride", 2, "FedStride", oneof="rule"),
         f("fed_rec", 3, "FedRec", oneof="rule"),
         f("pwa", 4, "PWA", oneof="rule"),
         f("fed_async", 6, "FedAsync", oneof="rule"),
         f("aggregation_rule_specs", 5, "AggregationRuleSpecs")),
    _Msg("AggregationRuleSpecs",
         _Enum("ScalingFactor", UNKNOWN=0, NUM_COMPLETED_BATCHES=1, NUM_PARTICIPANTS=2,
               NUM_TRAINING_EXAMPLES=3),
         f("scaling_factor", 1, "enum:AggregationRuleSpecs.ScalingFactor")),
    _Msg("FedAvg"),
    _Msg("FedStride", f("stride_length", 1, "uint32")),
    _Msg("FedRec"),
    _Msg("FedAsync",  # extension: C <- (1 - a_t) C + a_t x, a_t = alpha * s(t - tau)
         _Enum("StalenessFunction", CONSTANT=0, POLYNOMIAL=1, HINGE=2),
         f("alpha", 1, "float"),
         f("staleness_fn", 2, "enum:FedAsync.StalenessFunction"),
         f("a", 3, "float"),
         f("b", 4, "float")),
    _Msg("HESchemeConfig",
         f("enabled", 1, "bool"),
         f("crypto_context_file", 2, "string"),
         f("public_key_file", 3, "string"),
         f("private_key_file", 4, "string"),
         f("empty_scheme_config", 5, "EmptySchemeConfig", oneof="config"),
         f("ckks_scheme_config", 6, "CKKSSchemeConfig", oneof="config")),
    _Msg("EmptySchemeConfig"),
    _Msg("CKKSSchemeConfig", f("batch_size", 1, "uint32"), f("scaling_factor_bits", 2, "uint32")),
    _Msg("PWA", f("he_scheme_config", 1, "HESchemeConfig")),
    _Msg("GlobalModelSpecs", f("aggregation_rule", 1, "AggregationRule"),
         f("learners_participation_ratio", 2, "float")),
    _Msg("CommunicationSpecs",
         _Enum("Protocol", UNKNOWN=0, SYNCHRONOUS=1, ASYNCHRONOUS=2, SEMI_SYNCHRONOUS=3),
         f("protocol", 1, "enum:CommunicationSpecs.Protocol"),
         f("protocol_specs", 2, "ProtocolSpecs")),
    _Msg("ProtocolSpecs", f("semi_sync_lambda", 1, "int32"),
         f("semi_sync_recompute_num_updates", 2, "bool"),
         f("task_deadline_secs", 15, "float")),
    _Msg("LearnerDescriptor", f("id", 1, "string"), f("auth_token", 2, "string"),
         f("server_entity", 3, "ServerEntity"), f("dataset_spec", 4, "DatasetSpec")),
    _Msg("LearnerState", f("learner", 1, "LearnerDescriptor"), rep("model", 2, "Model")),
    _Msg("FederatedTaskRuntimeMetadata",
         f("global_iteration", 1, "uint32"),
         f("started_at", 2, TS),
         f("completed_at", 3, TS),
         rep("assigned_to_learner_id", 4, "string"),
         rep("completed_by_learner_id", 5, "string"),
         mp("train_task_submitted_at", 6, "string", TS),
         mp("train_task_received_at", 7, "string", TS),
         mp("eval_task_submitted_at", 8, "string", TS),
         mp("eval_task_received_at", 9, "string", TS),
         mp("model_insertion_duration_ms", 10, "string", "double"),
         mp("model_selection_duration_ms", 11, "string", "double"),
         f("model_aggregation_started_at", 12, TS),
         f("model_aggregation_completed_at", 13, TS),
         f("model_aggregation_total_duration_ms", 14, "double"),
         rep("model_aggregation_block_size", 15, "double"),
         rep("model_aggregation_block_memory_kb", 16, "double"),
         rep("model_aggregation_block_duration_ms", 17, "double"),
         rep("model_tensor_quantifiers", 18, "TensorQuantifier")),
]

# --------------------------------------------------------------------------
# controller.proto / learner.proto
# --------------------------------------------------------------------------
CONTROLLER = "metisfl/proto/controller.proto"
CONTROLLER_RPCS = [
    "GetCommunityModelEvaluationLineage", "GetCommunityModelLineage",
    "GetLearnerLocalModelLineage", "GetLocalTaskLineage", "GetRuntimeMetadataLineage",
    "GetParticipatingLearners", "GetServicesHealthStatus", "JoinFederation",
    "LeaveFederation", "MarkTaskCompleted", "ReplaceCommunityModel", "ShutDown",
]
_controller_msgs = [
    _Msg("GetCommunityModelEvaluationLineageRequest", f("num_backtracks", 1, "int32")),
    _Msg("GetCommunityModelEvaluationLineageResponse",
         rep("community_evaluation", 1, "CommunityModelEvaluation")),
    _Msg("GetCommunityModelLineageRequest", f("num_backtracks", 1, "int32")),
    _Msg("GetCommunityModelLineageResponse", rep("federated_models", 1, "FederatedModel")),
    _Msg("GetLocalTaskLineageRequest", f("num_backtracks", 1, "int32"),
         rep("learner_ids", 2, "string")),
    _Msg("GetLocalTaskLineageResponse", mp("learner_task", 1, "string", "LocalTasksMetadata")),
    _Msg("GetLearnerLocalModelLineageRequest", f("num_backtracks", 1, "int32"),
         rep("server_entity", 2, "ServerEntity")),
    _Msg("GetLearnerLocalModelLineageResponse",
         rep("learner_local_model", 1, "LearnerLocalModelResponse")),
    _Msg("GetRuntimeMetadataLineageRequest", f("num_backtracks", 1, "int32")),
    _Msg("GetRuntimeMetadataLineageResponse",
         rep("metadata", 1, "FederatedTaskRuntimeMetadata"), f("json_metadata", 2, "string")),
    _Msg("GetParticipatingLearnersRequest"),
    _Msg("GetParticipatingLearnersResponse", rep("learner", 1, "LearnerDescriptor")),
    _Msg("FabricMembership",  # extension: this learner is rank `rank` of a `world`-GPU peer-memory fabric
         f("fabric_id", 1, "string"), f("rank", 2, "uint32"), f("world", 3, "uint32"),
         f("staging_slots", 4, "uint32"),
         rep("endpoints", 5, "string"),             # "host:port" of every rank's LearnerService, by rank
         rep("public_certificates", 6, "bytes"),    # PEM streams by rank (empty without TLS)
         f("peer_mapped", 7, "bool")),              # ranks read each other's HBM; false = CPU tier (collectives)
    _Msg("JoinFederationRequest", f("server_entity", 1, "ServerEntity"),
         f("local_dataset_spec", 2, "DatasetSpec"), f("fabric", 15, "FabricMembership")),
    _Msg("JoinFederationResponse", f("ack", 1, "Ack"), f("learner_id", 2, "string"),
         f("auth_token", 3, "string"), f("ssl_config", 4, "SSLConfig")),
    _Msg("LearnerLocalModelResponse", f("server_entity", 1, "ServerEntity"),
         rep("model", 2, "Model")),
    _Msg("MarkTaskCompletedRequest", f("learner_id", 1, "string"), f("auth_token", 2, "string"),
         f("task", 3, "CompletedLearningTask")),
    _Msg("LearnerExecutionAuxMetadata", f("json_response", 1, "string")),
    _Msg("MarkTaskCompletedResponse", f("ack", 1, "Ack")),
    _Msg("LeaveFederationRequest", f("learner_id", 1, "string"), f("auth_token", 2, "string")),
    _Msg("LeaveFederationResponse", f("ack", 1, "Ack")),
    _Msg("ReplaceCommunityModelRequest", f("model", 1, "FederatedModel")),
    _Msg("ReplaceCommunityModelResponse", f("ack", 1, "Ack")),
]

LEARNER = "metisfl/proto/learner.proto"
LEARNER_RPCS = ["EvaluateModel", "GetServicesHealthStatus", "RunTask", "ShutDown", "AggregateOnDevice"]
_learner_msgs = [
    _Msg("EvaluateModelRequest",
         f("model", 1, "Model"),
         f("batch_size", 2, "uint32"),
         _Enum("dataset_to_eval", TRAINING=0, TEST=1, VALIDATION=2),
         rep("evaluation_dataset", 3, "enum:EvaluateModelRequest.dataset_to_eval"),
         f("metrics", 4, "EvaluationMetrics")),
    _Msg("EvaluateModelResponse", f("evaluations", 1, "ModelEvaluations")),
    _Msg("RunTaskRequest", f("federated_model", 1, "FederatedModel"),
         f("task", 2, "LearningTask"), f("hyperparameters", 3, "Hyperparameters")),
    _Msg("RunTaskResponse", f("ack", 1, "Ack")),
    # ---- extension: controller-commanded aggregation over device slots -------------------------
    _Msg("DeviceModelSource",
         f("learner_id", 1, "string"),
         f("device_rank", 2, "uint32"), f("slot", 3, "uint32"), f("version", 4, "uint64"),
         f("weight", 5, "double"),
         f("has_previous", 6, "bool"), f("previous_slot", 7, "uint32"), f("previous_weight", 8, "double"),
         f("staged_model", 9, "Model")),       # off-box learner: wire model, unpacked into a staging slot (K12)
    _Msg("AggregateOnDeviceRequest",
         f("rule", 1, "string"),               # FedAvg | FedStride | FedRec | FedAsync
         f("global_iteration", 2, "uint32"),
         rep("sources", 3, "DeviceModelSource"),
         f("mix", 4, "double"),                # FedAsync: alpha * s(staleness)
         f("z", 5, "double"),                  # FedRec: normaliser after this fold
         f("first_block", 6, "bool"), f("last_block", 7, "bool"),
         f("quantify", 8, "bool"),
         f("return_model_bytes", 9, "bool"),
         f("executor_rank", 10, "uint32")),
    _Msg("AggregateOnDeviceResponse",
         f("ack", 1, "Ack"),
         f("device_ms", 2, "double"),          # CUDA-event time of this rank's launch(es)
         rep("quantifiers", 3, "TensorQuantifier"),
         f("community_ref", 4, "Model"),       # control-only: metadata + device_slot references
         f("community_model", 5, "Model"),     # with bytes, only when asked for
         f("version", 6, "uint64")),
]

_TSFILE = "google/protobuf/timestamp.proto"
FILES = [
    _build_file(MODEL, [], _model_msgs),
    _build_file(COMMON, [_TSFILE], _common_msgs),
    _build_file(METIS, [MODEL, _TSFILE], _metis_msgs),
    _build_file(CONTROLLER, [METIS, MODEL, COMMON], _controller_msgs,
                services=[("ControllerService", CONTROLLER_RPCS)]),
    _build_file(LEARNER, [METIS, MODEL, COMMON], _learner_msgs,
                services=[("LearnerService", LEARNER_RPCS)]),
]

# Register in the default pool so Timestamp resolves and json_format/text_format work.
POOL = descriptor_pool.Default()
_FILE_DESCRIPTORS = {}
for _fdp in FILES:
    try:
        _FILE_DESCRIPTORS[_fdp.name] = POOL.Add(_fdp) or POOL.FindFileByName(_fdp.name)
    except TypeError:  # pragma: no cover - very old runtimes
        POOL.AddSerializedFile(_fdp.SerializeToString())
        _FILE_DESCRIPTORS[_fdp.name] = POOL.FindFileByName(_fdp.name)
    if _FILE_DESCRIPTORS[_fdp.name] is None:
        _FILE_DESCRIPTORS[_fdp.name] = POOL.FindFileByName(_fdp.name)


def file_descriptor(name):
    return _FILE_DESCRIPTORS[name]


def message_class(name: str):
    """Concrete class for ``metisfl.<name>`` (dotted path for nested types)."""
    return message_factory.GetMessageClass(POOL.FindMessageTypeByName(f"{PKG}.{name}"))


def export_messages(file_name: str, namespace: dict):
    """Populate a ``*_pb2`` module namespace with the file's top-level message classes."""
    fd = _FILE_DESCRIPTORS[file_name]
    namespace["DESCRIPTOR"] = fd
    for mname in fd.message_types_by_name:
        namespace[mname] = message_class(mname)


def to_proto_text() -> str:
    """Render the schema as .proto-like text (documentation only)."""
    out = []
    for fdp in FILES:
        out.append(f"// ---- {fdp.name}\n{fdp}")
    return "\n".join(out)
