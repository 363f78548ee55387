"""CPU tests (no GPU): shrink the part of the oracle that rests on recollection of prost 0.6.0 alone.

The reference has no tests or vectors for this path and prost cannot be built here (no rustc, crates not vendored), so
the oracle (oracle/cpu_ref.cpp) is compared with an independent protobuf implementation - upb, through message classes
rebuilt at run time from the reference's own IDL (protos/tensorflow_serving/apis/predict.proto:12-40, model.proto:9-33;
tests/golden/make_golden.py holds the descriptors) - on WHOLE PredictRequest / PredictResponse / ModelSpec messages:

  * encode: random messages (hypothesis), outside the cases where prost and upb are known to differ (SURVEY.md Appendix E:
    empty map key, default map value, version_label ordering), must serialise to the same bytes;
  * decode: upb's bytes decode to the same values;
  * accept / reject: a corpus of byte-level mutations of valid messages must be accepted or rejected alike, except for the
    differences enumerated in PROST_UPB_DISAGREEMENTS below (each with the prost rule behind it).
"""
import os
import sys

import numpy as np
import pytest

import tfs_b200 as tfs
from tfs_b200 import abi, workloads

import helpers
import oracle_bind as ob

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
try:
    import make_golden as mg
    from google.protobuf.message import DecodeError
except Exception:  # pragma: no cover - tensorboard's compiled TensorProto is part of the image
    mg = None

pytestmark = pytest.mark.skipif(mg is None, reason="tensorboard / protobuf (upb) not importable")


def _tensor_desc(draw_ints, rng_bytes, dtype, shape, i64, i32, f32bits, bools, content):
    fields = {}
    if i64:
        fields["int64_val"] = mg.arr_hex(np.array(i64, "<i8"))
    if i32:
        fields["int_val"] = mg.arr_hex(np.array(i32, "<i4"))
    if f32bits:
        fields["float_val"] = mg.arr_hex(np.array(f32bits, "<u4"))
    if bools:
        fields["bool_val"] = mg.arr_hex(np.array(bools, "u1"))
    if content:
        fields["tensor_content"] = content.hex()
    return {"dtype": dtype, "shape": shape, "unknown_rank": False, "version_number": 0, "fields": fields}


def test_whole_message_differential_vs_upb():
    from hypothesis import given, settings, strategies as st

    ints64 = st.one_of(st.integers(-(1 << 63), (1 << 63) - 1), st.integers(-130, 130),
                       st.sampled_from([127, 128, 16383, 16384, -1, (1 << 63) - 1, -(1 << 63)]))
    # no signalling NaNs: python float access quiets them inside upb (last row)
    f32 = st.integers(0, (1 << 32) - 1).filter(lambda b: not ((b & 0x7F800000) == 0x7F800000 and (b & 0x007FFFFF) != 0 and not (b & 0x00400000)))
    tensor = st.fixed_dictionaries({
        "dtype": st.integers(1, 23),  # a tensor that is all defaults is an Appendix E case (value sub-field omitted by prost)
        "shape": st.one_of(st.none(), st.lists(st.integers(-1, 1 << 40), max_size=8)),
        "i64": st.lists(ints64, max_size=30), "i32": st.lists(st.integers(-(1 << 31), (1 << 31) - 1), max_size=30),
        "f32bits": st.lists(f32, max_size=30), "bools": st.lists(st.booleans(), max_size=10), "content": st.binary(max_size=40)})
    name = st.text(min_size=1, max_size=12)  # an empty map key is an Appendix E case
    spec = st.one_of(st.none(), st.fixed_dictionaries({
        "name": st.text(max_size=20), "signature_name": st.text(max_size=20),
        "version": st.one_of(st.none(), st.integers(-(1 << 63), (1 << 63) - 1))}))  # version_label: Appendix E

    @settings(max_examples=120, deadline=None)
    @given(kind=st.sampled_from(["request", "response"]), spec=spec,
           entries=st.lists(st.tuples(name, tensor), max_size=4, unique_by=lambda e: e[0]),
           filters=st.lists(st.text(max_size=10), max_size=3))
    def run(kind, spec, entries, filters):
        ents = [[k, _tensor_desc(None, None, **t)] for k, t in entries]
        desc = {"model_spec": spec, "entries": ents}
        if kind == "request":
            desc["output_filter"] = filters
        want = mg.upb_serialize({"kind": kind, "msg": desc})
        # a map has no order of its own (prost: HashMap iteration order; upb: its own sort): feed the oracle the entries in
        # the order upb happened to write them, found by reading upb's bytes back
        st0, back = ob.decode_bytes([want], helpers.KIND_OF[kind])
        assert st0 == 0
        order = [k for k, _ in (back[0].message.inputs if kind == "request" else back[0].message.outputs)]
        assert sorted(order) == sorted(k for k, _ in ents)
        desc["entries"] = sorted(ents, key=lambda e: order.index(e[0]))
        msg = helpers.message_from_desc(kind, desc)
        got = ob.encode_messages([msg], helpers.KIND_OF[kind])[0]
        assert got == want, (got.hex(), want.hex())
        st_, msgs = ob.decode_bytes([want], helpers.KIND_OF[kind])
        assert st_ == 0 and msgs[0].status == 0
        got_desc, want_desc = helpers.message_to_desc(kind, msgs[0].message), helpers.norm_spec(mg.upb_parse(kind, want))
        got_desc["entries"].sort(key=lambda e: e[0])
        want_desc["entries"].sort(key=lambda e: e[0])
        assert got_desc == want_desc
        # prost's encoded_len of the whole message
        assert ob.encode_size(tfs.build_batch([msg], helpers.KIND_OF[kind])) == [len(want)]

    run()


def test_model_spec_alone_vs_upb():
    # ModelSpec inside a request with nothing else: every combination of name / version / signature the reference can
    # build (src/lib.rs:433-447: the label is never set)
    for name in ("", "resnet", "m" * 200):
        for sig in ("", "serving_default"):
            for ver in (None, 0, 1, 123, -1, (1 << 63) - 1):
                desc = {"model_spec": {"name": name, "signature_name": sig, "version": ver}, "entries": [], "output_filter": []}
                want = mg.upb_serialize({"kind": "request", "msg": desc})
                got = ob.encode_messages([helpers.message_from_desc("request", desc)], abi.MSG_PREDICT_REQUEST)[0]
                assert got == want, (name, sig, ver, got.hex(), want.hex())


# Where prost 0.6.0 and upb are KNOWN to decide differently on malformed or odd input; everything else must agree.
#   key: (oracle accepts, upb accepts, oracle status)
PROST_UPB_DISAGREEMENTS = {
    (False, True, abi.TFS_ERR_WIRE_TYPE):
        "prost's derived merge_field checks the wire type of every KNOWN field and fails with 'invalid wire type: X (expected Y)'; "
        "upb files a known field number that arrives with the wrong wire type under unknown fields and goes on",
    (False, True, abi.TFS_ERR_UNSUPPORTED):
        "not a prost rule: valid wire that uses a field this library hands back to prost (Dim.name, rank > 8)",
    (True, False, abi.TFS_OK):
        "prost's hash_map::merge does not look at the wire type of the map field itself (SURVEY.md A.5) and accepts a 10th varint "
        "byte with high bits set; only reachable through the hand-written cases in test_oracle.py, never seen in this corpus",
}


def mutation_corpus(kind, n_cases, seed=99):
    return helpers.mutation_corpus(kind, n_cases, seed)


@pytest.mark.parametrize("kind", ["response", "request"])
def test_accept_reject_agreement_with_upb_on_the_mutation_corpus(kind):
    blobs = helpers.mutation_corpus(kind, 400)
    st_, want = ob.decode_bytes(blobs, helpers.KIND_OF[kind])
    cls = mg.PredictResponsePB if kind == "response" else mg.PredictRequestPB
    seen = {}
    for i, (b, w) in enumerate(zip(blobs, want)):
        pb = cls()
        try:
            pb.ParseFromString(b)
            upb_ok = True
        except DecodeError:
            upb_ok = False
        ours_ok = w.status == 0
        if ours_ok == upb_ok:
            if ours_ok:  # both accept: the values agree too (unknown fields aside, which prost drops)
                got = helpers.message_to_desc(kind, w.message)
                theirs = helpers.norm_spec(mg.upb_parse(kind, b))
                if got != theirs:
                    # the one value-level difference: a map entry with an UNKNOWN sub-field.  prost skips the sub-field and
                    # inserts the entry (hash_map::merge -> skip_field); upb keeps the whole entry as an unknown field of
                    # the parent.  The entries upb lacks must be exactly that many unknown fields under the map's number.
                    from google.protobuf.unknown_fields import UnknownFieldSet
                    map_field = 1 if kind == "response" else 2
                    n_unknown = sum(1 for f in UnknownFieldSet(pb) if f.field_number == map_field)
                    missing = [e for e in got["entries"] if e[0] not in {k for k, _ in theirs["entries"]}]
                    assert missing and 0 < len(missing) <= n_unknown, f"case {i}: {b.hex()}"
                    got = dict(got)
                    got["entries"] = [e for e in got["entries"] if e not in missing]
                    assert got == theirs, f"case {i}: {b.hex()}"
                    seen["entry with unknown sub-field"] = seen.get("entry with unknown sub-field", 0) + 1
            continue
        key = (ours_ok, upb_ok, w.status)
        assert key in PROST_UPB_DISAGREEMENTS, f"case {i}: oracle {abi.STATUS_NAMES.get(w.status)} vs upb accept={upb_ok}: {b.hex()}"
        seen[key] = seen.get(key, 0) + 1
    # the corpus does exercise the table's main row, and nothing outside it (asserted above)
    assert seen.get((False, True, abi.TFS_ERR_WIRE_TYPE), 0) > 0
    assert (True, False, abi.TFS_OK) not in seen
