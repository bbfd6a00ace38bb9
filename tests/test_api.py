"""JSON control API: codecs, validation, registry, bucketing (FlinkNetwork.scala:48-240), merging."""
import json

from omldm_b200.api.bucketing import MAX_PARAM_BUCKET_SIZE, bucket_response, reassemble, split
from omldm_b200.api.pojos import (DataInstance, JobStatistics, LearnerPOJO, Prediction, QueryResponse, Request,
                                  Statistics)
from omldm_b200.api.registry import PipelineMap
from omldm_b200.api.response_merger import ResponseConstructor
from omldm_b200.messages import HubMessage, NodeId, NodeType, RemoteCallIdentifier, SpokeMessage
from omldm_b200.stream.parsers import parse_data_instance, parse_request
from omldm_b200.utils.partitioners import hub_key, hub_message_partition, identity_partitioner

CREATE = {"id": 1, "request": "Create", "learner": {"name": "PA", "hyperParameters": {"C": 0.1}},
          "preProcessors": [{"name": "StandardScaler"}], "trainingConfiguration": {"protocol": "FGM"}}


def test_request_roundtrip_and_validation():
    r = Request.from_json(json.dumps(CREATE))
    assert r.is_valid() and r.learner.name == "PA" and r.preProcessors[0].name == "StandardScaler"
    assert Request.from_dict(r.to_dict()).to_dict() == r.to_dict()
    assert not Request.from_dict({"id": 1, "request": "Bogus"}).is_valid()
    assert not Request.from_dict({"id": 1, "request": "Create"}).is_valid()          # no learner
    assert not Request.from_dict({"id": 1, "request": "Query"}).is_valid()           # no requestId
    assert parse_request({"id": -3, "request": "Delete"}) is None


def test_data_instance_parser_drops_garbage_silently():
    ok = parse_data_instance('{"numericalFeatures":[1,2],"discreteFeatures":[3],"target":1.0,"operation":"training"}')
    assert ok.numericalFeatures == [1, 2] and ok.discreteFeatures == [3]
    assert parse_data_instance("EOS") is None
    assert parse_data_instance("{not json") is None
    assert parse_data_instance('{"numericalFeatures":[1],"operation":"other"}') is None


def test_pipeline_map_whitelist_and_routing():
    pm = PipelineMap(4)
    assert len(pm.flat_map(Request.from_dict(CREATE))) == 4                        # broadcast to every worker
    assert pm.flat_map(Request.from_dict(CREATE)) == []                            # duplicate id
    bad = dict(CREATE, id=2, learner={"name": "XGBoost"})
    assert pm.flat_map(Request.from_dict(bad)) == []
    badpp = dict(CREATE, id=3, preProcessors=[{"name": "PCA"}])
    assert pm.flat_map(Request.from_dict(badpp)) == []
    assert len(pm.flat_map(Request.from_dict({"id": 1, "request": "Query", "requestId": 9}))) == 4
    assert len(pm.flat_map(Request.from_dict({"id": 1, "request": "Update"}))) == 4
    ht = dict(CREATE, id=5, learner={"name": "HT"})
    pm.flat_map(Request.from_dict(ht))
    q = pm.flat_map(Request.from_dict({"id": 5, "request": "Query", "requestId": 1}))
    assert len(q) == 1 and q[0].destination.node_id == 0                           # HT: worker 0 only
    assert len(pm.flat_map(Request.from_dict({"id": 1, "request": "Delete"}))) == 4
    assert pm.flat_map(Request.from_dict({"id": 1, "request": "Query", "requestId": 2})) == []
    pm2 = PipelineMap(4)
    pm2.load_state_dict(pm.state_dict())
    assert set(pm2.nodes) == set(pm.nodes)


def test_split_bucket_keys_and_scalars():
    b = split({"a": list(range(25000)), "b": 0.5, "c": [7]})
    assert sorted(b) == [0, 1, 2]
    assert set(b[0]) == {"a[0-9999]", "b", "c"} and b[0]["c"] == 7 and b[0]["b"] == 0.5
    assert list(b[1]) == ["a[10000-19999]"] and list(b[2]) == ["a[20000-24999]"]
    assert len(b[2]["a[20000-24999]"]) == 5000
    exact = split({"a": list(range(2 * MAX_PARAM_BUCKET_SIZE))})                  # reference bug fixed: no empty slice
    assert list(exact[1]) == ["a[10000-19999]"] and len(exact[1]["a[10000-19999]"]) == 10000


def test_bucket_response_layout_and_reassembly():
    resp = QueryResponse(5, None, 1, [], LearnerPOJO("PA", {"C": 0.1}, {"a": list(range(35000)), "b": 0.5}, None),
                         "Synchronous", 10, 0.1, 0.2, 0.3)
    frags = bucket_response(resp)
    assert len(frags) == 4 and [f.bucketIndex for f in frags] == [0, 1, 2, 3]
    assert frags[0].dataFitted is None and frags[0].protocol is None               # bucket 0: learner maps only
    assert frags[3].dataFitted == 10 and frags[3].protocol == "Synchronous"        # last parameter bucket: scalars
    assert all(f.loss is None for f in frags[:3])
    whole = reassemble(frags)
    assert whole.learner.parameters["a"] == list(range(35000)) and whole.score == 0.3
    small = QueryResponse(6, None, 1, [], LearnerPOJO("PA", None, {"a": list(range(15000))}, None), "x", 1, 0., 0., 0.)
    assert len(bucket_response(small)) == 1                                        # <= 2 buckets: sent whole
    probe = QueryResponse(-1, None, 1, None, LearnerPOJO("PA", None, {"a": list(range(50000))}, None))
    assert bucket_response(probe) == [probe]


def test_response_constructor_sum_and_mean():
    rc = ResponseConstructor(3)
    out = []
    for i in range(3):
        out += rc.flat_map(QueryResponse(7, None, 1, None, LearnerPOJO("PA"), "Asynchronous", 10 * (i + 1),
                                         float(i), float(2 * i), 0.5 + 0.1 * i))
    assert len(out) == 1
    m = out[0]
    assert m.dataFitted == 60 and abs(m.loss - 1.0) < 1e-12 and abs(m.cumulativeLoss - 2.0) < 1e-12
    assert abs(m.score - 0.6) < 1e-12 and m.protocol == "Asynchronous"
    frag = QueryResponse(8, 0, 1, None, LearnerPOJO("PA"))                          # bucket fragment passes through
    assert rc.flat_map(frag) == [frag]
    rc.expect(9, 1)
    assert len(rc.flat_map(QueryResponse(9, None, 1, None, LearnerPOJO("HT"), "SingleLearner", 5, 0., 0., 1.))) == 1


def test_messages_sizes_fanout_partitioners():
    hm = HubMessage(3, [RemoteCallIdentifier("update_model")] * 2, NodeId(NodeType.HUB, 0),
                    [NodeId(NodeType.SPOKE, 0), NodeId(NodeType.SPOKE, 5)], [1.0, 2.0, 3.0], None)
    cms = hm.fan_out()
    assert [c.destination.node_id for c in cms] == [0, 5] and cms[0].data is cms[1].data
    assert hm.get_size() > 4 and SpokeMessage().get_size() == 4
    assert HubMessage().fan_out() == []                                             # termination message
    assert hub_message_partition(HubMessage(), 36) == 0
    assert hub_message_partition(hm, 36) == 3
    single = HubMessage(3, [RemoteCallIdentifier("x")], None, [NodeId(NodeType.SPOKE, 40)])
    assert hub_message_partition(single, 36) == 4
    assert identity_partitioner(5, 8) == 5 and hub_key(3, 1) == "3_1"


def test_statistics_and_prediction_json():
    st = Statistics(1, "FGM")
    st.update_stats(2, 100, 3); st.update_fitted(10); st.update_score(0.5); st.update_mean_buffer_size(1.5)
    js = JobStatistics("job", 4, 12.5, [st])
    d = json.loads(str(js))
    assert d["parallelism"] == 4 and d["statistics"][0]["bytes"] == 100 and d["statistics"][0]["blocks"] == 3
    p = Prediction(1, DataInstance([1.0, 2.0], operation="forecasting"), 0.25)
    assert json.loads(str(p))["prediction"] == 0.25


def test_compat_api_surface():
    """Names of the out-of-tree API the reference touches: VectorBias, Parsing, CountableSerial, Mergeable."""
    import numpy as np
    from omldm_b200.api.compat import (CountableSerial, Mergeable, VectorBias, double_parsing, integer_parsing)
    from omldm_b200.api.pojos import Request
    from omldm_b200.oracle.learners import make_oracle_learner
    from omldm_b200.parallel.network import DataSet
    pa = make_oracle_learner("PA", {"C": 0.1})
    pa.ensure_init(3)
    pa.w[:] = [1.0, -2.0, 0.5, 0.25]
    vb = VectorBias.from_flat(pa.w)
    assert vb.bias == 0.25 and vb.to_parameters() == pa.to_parameters() and vb == VectorBias.from_parameters(pa.to_parameters())
    assert vb.predict([1, 1, 2]) == 1.0 - 2.0 + 1.0 + 0.25 and vb.get_size() == 32 and len(vb) == 4
    avg = 0.5 * (vb + VectorBias([3.0, 0.0, 0.5], 0.75))
    assert np.allclose(avg.flat(), [2.0, -1.0, 0.5, 0.5])
    assert integer_parsing({"HubParallelism": "3"}, "HubParallelism", 1) == 3
    assert integer_parsing({"HubParallelism": "x"}, "HubParallelism", 1) == 1 and integer_parsing(None, "k", 7) == 7
    assert double_parsing({"t": "0.5"}, "t", 1.0) == 0.5 and double_parsing({}, "t", 1.0) == 1.0
    req = Request.from_dict({"id": 1, "request": "Create", "learner": {"name": "PA"}})
    assert isinstance(req, CountableSerial) and isinstance(vb, CountableSerial)
    assert isinstance(DataSet(4), Mergeable)
    # the protocols read their configuration through the same rule: a malformed value falls back to the default
    from omldm_b200.parallel.host_protocols import cfg_float, cfg_int
    assert cfg_int({"staleness": "4"}, ["staleness", "s"], 3) == 4 and cfg_int({"s": "many"}, ["staleness", "s"], 3) == 3
    assert cfg_float({"threshold": None, "T": "0.25"}, ["threshold", "T"], 0.5) == 0.25 and cfg_float(None, ["T"], 0.5) == 0.5


def test_math_vector_and_point_types():
    """`mlAPI.math` names registered by the reference (CommonUtils.scala:17-23) and built by DataPointParser.scala:20-46."""
    import numpy as np
    import pytest
    from omldm_b200.stream.points import (DenseVector, ForecastingPoint, LabeledPoint, SparseVector, TrainingPoint,
                                          UnlabeledPoint)
    s = SparseVector(10, [3, 1, 3], [1.0, 2.0, 0.5])             # duplicate indices accumulate
    d = DenseVector(np.arange(10))
    assert s.indices.tolist() == [1, 3] and s.values.tolist() == [2.0, 1.5]
    assert s.dot(d) == d.dot(s) == 2.0 * 1 + 1.5 * 3 and s.to_dense() == s and len(s) == 10
    assert s.dot(SparseVector(10, [3, 7], [2.0, 9.0])) == 3.0
    idx, val = s.to_ell(4)
    assert idx.dtype == np.int32 and idx.tolist() == [1, 3, 0, 0] and val.tolist() == [2.0, 1.5, 0.0, 0.0]
    with pytest.raises(ValueError):
        SparseVector(3, [5], [1.0])
    lp = LabeledPoint(1.0, s, [1, 2], ["a"], "raw")
    assert isinstance(lp, TrainingPoint) and lp.target == 1.0 and lp.features().shape == (12,)
    up = lp.as_unlabeled_point()
    assert isinstance(up, UnlabeledPoint) and isinstance(up, ForecastingPoint) and up.raw == "raw"
    from omldm_b200.stream.points import LearningPoint, Point, Vector
    assert isinstance(lp, Point) and isinstance(up, LearningPoint) and isinstance(s, Vector) and isinstance(d, Vector)


def test_parameter_server_base_names():
    """`MLParameterServer` accessors the hub reads (FlinkHub.scala:98-127) and the `NodeInstance` base (NodeGenerator.scala:8-9)."""
    from omldm_b200.parallel.host_protocols import (AsynchronousWorker, MLParameterServer, NodeInstance,
                                                    SynchronousParameterServer)
    ps = SynchronousParameterServer()
    assert isinstance(ps, MLParameterServer) and isinstance(ps, NodeInstance)
    ps.set_idx(3)
    assert ps.get_idx() == 3 and ps.get_node_id() is None and ps.get_number_of_fitted_data() == 0
    assert ps.get_protocol_statistics().protocol == "Synchronous" and ps.get_learning_curve() == []
    assert issubclass(AsynchronousWorker, NodeInstance)
