"""JSON parsers for the request and data streams.

Reference: utils/parsers/RequestParser.scala:12-16 (Jackson -> Request, keep if isValid; a parse
error there fails the job - here it raises ValueError, which the CLI treats as fatal),
utils/parsers/DataInstanceParser.scala:12-22 (skips the literal "EOS", swallows all errors),
utils/parsers/StringToArrayDoublesParser.scala (CSV line -> doubles), and the unused Kafka
deserializers utils/deserializers/{DataInstance,Request}Deserializer.scala:25-32 that attach
(topic, partition, key, offset, timestamp) metadata; utils/serializers/GenericSerializer.scala
(object -> JSON bytes).
"""
from __future__ import annotations

import json
from typing import Any, List, Optional, Union

from ..api.pojos import DataInstance, Request


def parse_request(rec: Union[str, bytes, dict, Request]) -> Optional[Request]:
    if isinstance(rec, Request):
        r = rec
    else:
        try:
            d = rec if isinstance(rec, dict) else json.loads(rec)
        except (json.JSONDecodeError, TypeError) as e:
            raise ValueError(f"malformed request: {e}") from e
        r = Request.from_dict(d)
    return r if r.is_valid() else None


def parse_data_instance(rec: Union[str, bytes, dict, DataInstance]) -> Optional[DataInstance]:
    if isinstance(rec, DataInstance):
        return rec if rec.is_valid() else None
    try:
        if isinstance(rec, (str, bytes)):
            s = rec.decode() if isinstance(rec, bytes) else rec
            if s.strip() == "EOS":
                return None
            d = json.loads(s)
        else:
            d = rec
        di = DataInstance.from_dict(d)
        return di if di.is_valid() else None
    except Exception:
        return None   # malformed data records are dropped silently


def parse_doubles(line: str, sep: str = ",") -> List[float]:
    return [float(t) for t in line.strip().split(sep) if t.strip()]


def deserialize_with_metadata(kind: str, payload: bytes, topic: str, partition: int, key: Any, offset: int,
                              timestamp: int):
    obj = parse_request(payload) if kind == "request" else parse_data_instance(payload)
    if obj is not None:
        obj.set_metadata(topic, partition, key, offset, timestamp)
    return obj


def generic_serialize(obj: Any) -> bytes:
    if hasattr(obj, "to_json"):
        return obj.to_json().encode()
    return json.dumps(obj).encode()
