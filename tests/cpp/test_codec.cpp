// GPU test of the C++ host mirror (include/tfs_codec.hpp): GpuCodec::encoder()/decoder() against the known-answer
// vectors of SURVEY.md Appendix B and against the CPU oracle (checker only), through the C ABI.
// Built by __graft_entry__.build(); run by tests/test_gpu_cpp_codec.py on the GPU box.
#include <cstdio>
#include <cstdlib>
#include <random>
#include <string>

#include "../../include/tfs_codec.hpp"

extern "C" int32_t oracle_encode(const tfs_batch_t* b, uint8_t* dst, uint64_t cap, uint32_t flags, uint64_t* off, uint64_t* len, uint64_t* total);

using namespace tfs_codec;

static int failures = 0;
#define CHECK(cond)                                                       \
  do {                                                                    \
    if (!(cond)) {                                                        \
      std::printf("CHECK FAILED %s:%d: %s\n", __FILE__, __LINE__, #cond); \
      ++failures;                                                         \
    }                                                                     \
  } while (0)

static std::string hex(const std::vector<uint8_t>& v) {
  static const char* d = "0123456789abcdef";
  std::string s;
  for (uint8_t b : v) {
    s.push_back(d[b >> 4]);
    s.push_back(d[b & 15]);
  }
  return s;
}

static std::vector<uint8_t> oracle_bytes(const PredictRequest& r) {
  detail::Described d;
  d.add_message(TFS_MSG_PREDICT_REQUEST, r.model_spec, r.inputs, &r.output_filter);
  const tfs_batch_t b = d.batch();
  std::vector<uint8_t> out(tfs_encode_bound(&b, 0) + 64);
  uint64_t off = 0, len = 0, total = 0;
  if (oracle_encode(&b, out.data(), out.size(), 0, &off, &len, &total) != 0) std::abort();
  out.resize(total);
  return out;
}

int main() {
  GpuCodec codec;  // throws without a CUDA device: no CPU fallback
  GpuEncoder enc = codec.encoder();
  GpuDecoder dec = codec.decoder();

  // Appendix B.1: the half_plus_two request
  PredictRequest req;
  req.model_spec = ModelSpec{"half_plus_two", std::nullopt, std::nullopt, "serving_default"};
  TensorProto x;
  x.dtype = 1;
  x.tensor_shape = TensorShapeProto{{{3, ""}}, false};
  x.float_val = {1.0f, 2.0f, 5.0f};
  req.inputs.emplace_back("x", x);
  std::vector<uint8_t> wire;
  enc.encode(req, wire);
  CHECK(hex(wire) == "0a200a0d68616c665f706c75735f74776f1a0f73657276696e675f64656661756c74121b0a0178121608011204120208032a0c0000803f000000400000a040");
  CHECK(enc.encoded_len(req) == 63);

  // Appendix B.2: the response, decoded
  std::vector<uint8_t> resp_wire;
  {
    const char* h = "0a1b0a0179121608011204120208032a0c00002040000040400000904012240a0d68616c665f706c75735f74776f1202087b1a0f73657276696e675f64656661756c74";
    for (size_t i = 0; h[i]; i += 2) resp_wire.push_back(static_cast<uint8_t>(std::stoi(std::string(h + i, 2), nullptr, 16)));
  }
  std::vector<uint8_t> body = resp_wire;
  auto resp = dec.decode(body);
  CHECK(resp.has_value() && body.empty());
  CHECK(resp->outputs.size() == 1 && resp->outputs[0].first == "y");
  CHECK(resp->outputs[0].second.float_val == std::vector<float>({2.5f, 3.0f, 4.5f}));
  CHECK(resp->outputs[0].second.dtype == 1 && resp->outputs[0].second.tensor_shape->dim.size() == 1 && resp->outputs[0].second.tensor_shape->dim[0].size == 3);
  CHECK(resp->model_spec && resp->model_spec->name == "half_plus_two" && resp->model_spec->version == 123 && resp->model_spec->signature_name == "serving_default");
  // and re-encoded it is the same bytes again
  std::vector<uint8_t> again;
  enc.encode(*resp, again);
  CHECK(again == resp_wire);

  // a BERT-shaped request with every payload kind, against the oracle, then round-tripped
  std::mt19937_64 rng(3001);
  PredictRequest big;
  big.model_spec = ModelSpec{"bert", 7, std::nullopt, "serving_default"};
  big.output_filter = {"logits", ""};
  TensorProto ids, mix;
  ids.dtype = 9;
  ids.tensor_shape = TensorShapeProto{{{16, ""}, {512, ""}}, false};
  for (int i = 0; i < 16 * 512; ++i) ids.int64_val.push_back(static_cast<int64_t>(rng() % 30522) - (i % 97 == 0 ? 40000 : 0));
  mix.dtype = 1;
  for (int i = 0; i < 5000; ++i) {
    mix.float_val.push_back(static_cast<float>(rng() % 1000) / 7.0f);
    mix.double_val.push_back(static_cast<double>(rng() % 1000) / 3.0);
    mix.int_val.push_back(static_cast<int32_t>(rng()));
    mix.half_val.push_back(static_cast<int32_t>(rng() & 0xFFFF));
    mix.bool_val.push_back(static_cast<uint8_t>(rng() & 1));
    mix.uint32_val.push_back(static_cast<uint32_t>(rng()));
    mix.uint64_val.push_back(rng() >> (rng() % 64));
    mix.tensor_content.push_back(static_cast<uint8_t>(rng()));
  }
  for (int i = 0; i < 300; ++i) {  // string_val: Vec<Vec<u8>>, sizes 0 .. 70 000
    std::vector<uint8_t> e(i % 7 == 0 ? 0 : (i == 150 ? 70000 : rng() % 60));
    for (auto& b : e) b = static_cast<uint8_t>(rng());
    mix.string_val.push_back(std::move(e));
  }
  big.inputs.emplace_back("input_ids", ids);
  big.inputs.emplace_back("mix", mix);
  std::vector<uint8_t> w2;
  enc.encode(big, w2);
  CHECK(w2 == oracle_bytes(big));
  std::vector<uint8_t> b2 = w2;
  auto back = dec.decode_request(b2);
  CHECK(back && back->inputs.size() == 2 && back->inputs[0].first == "input_ids");
  CHECK(back->inputs[0].second.int64_val == ids.int64_val);
  CHECK(back->inputs[1].second.float_val == mix.float_val && back->inputs[1].second.double_val == mix.double_val);
  CHECK(back->inputs[1].second.int_val == mix.int_val && back->inputs[1].second.half_val == mix.half_val);
  CHECK(back->inputs[1].second.bool_val == mix.bool_val && back->inputs[1].second.uint32_val == mix.uint32_val);
  CHECK(back->inputs[1].second.uint64_val == mix.uint64_val && back->inputs[1].second.tensor_content == mix.tensor_content);
  CHECK(back->inputs[1].second.string_val == mix.string_val);
  CHECK(back->output_filter == big.output_filter && back->model_spec->version == 7);

  // a batch of requests in one launch; each message's bytes equal the single-message encoding
  std::vector<PredictRequest> window = {req, big, req};
  std::vector<uint8_t> w3;
  std::vector<std::pair<uint64_t, uint64_t>> spans;
  enc.encode_batch(window.data(), window.size(), w3, &spans);
  CHECK(spans.size() == 3);
  CHECK(std::vector<uint8_t>(w3.begin() + spans[1].first, w3.begin() + spans[1].first + spans[1].second) == w2);
  CHECK(std::vector<uint8_t>(w3.begin() + spans[2].first, w3.begin() + spans[2].first + spans[2].second) == wire);

  // Classify / Regress input: build_input(HashMap<name, Payload>) -> Input, known-answer bytes (hand-checked, tests/test_oracle.py)
  {
    Features fm;
    fm.emplace_back("age", Payload::Of(std::vector<int64_t>{29}));
    fm.emplace_back("w", Payload::Of(std::vector<float>{1.5f, 2.5f}));
    fm.emplace_back("name", Payload::Of(std::vector<std::vector<uint8_t>>{{'b', 'o', 'b'}, {}}));
    Input in = build_input(fm);
    in.examples.push_back(Example{});
    Features third;
    third.emplace_back("", Feature{});
    third.emplace_back("e", Payload::Of(std::vector<float>{}));
    in.examples.push_back(Example{third});
    std::vector<uint8_t> wi;
    enc.encode(in, wi);
    const std::string want =
        "0a4b0a360a340a0c0a0361676512051a030a011d0a110a0177120c120a0a080000c03f000020400a110a046e616d6512090a070a03626f620a000a020a000a0d0a0b0a000a070a016512021200";
    CHECK(hex(wi) == want);
    ClassificationRequest cr;
    cr.model_spec = ModelSpec{"m", std::nullopt, std::nullopt, "s"};
    cr.input = in;
    std::vector<uint8_t> wc;
    enc.encode(cr, wc);
    CHECK(hex(wc) == "0a060a016d1a0173124d" + want);
  }

  // malformed input surfaces as a Status carrying prost's DecodeError class
  std::vector<uint8_t> bad = {0x0a, 0x10, 0x0a, 0x01, 0x6b};
  try {
    dec.decode(bad);
    CHECK(false);
  } catch (const Status& s) {
    CHECK(s.code == TFS_ERR_BUFFER_UNDERFLOW && std::string(s.what()) == "buffer underflow");
  }
  // a duplicated map key: the last entry wins (HashMap::insert)
  std::vector<uint8_t> dup = {0x0a, 0x07, 0x0a, 0x01, 0x6b, 0x12, 0x02, 0x08, 0x01, 0x0a, 0x07, 0x0a, 0x01, 0x6b, 0x12, 0x02, 0x08, 0x03};
  auto d2 = dec.decode(dup);
  CHECK(d2 && d2->outputs.size() == 1 && d2->outputs[0].second.dtype == 3);

  // the metadata call (src/lib.rs:300-312): the request the reference builds, byte for byte (upb and the oracle agree on
  // these bytes, tests/test_metadata_messages.py), and a response with one signature_def entry
  GetModelMetadataRequest mreq;
  mreq.model_spec = ModelSpec{"resnet", 3, std::nullopt, ""};
  mreq.metadata_field = {"signature_def"};
  std::vector<uint8_t> mw;
  enc.encode(mreq, mw);
  CHECK(hex(mw) == "0a0c0a067265736e657412020803120d7369676e61747572655f646566");
  std::vector<uint8_t> mresp = {0x0a, 0x03, 0x0a, 0x01, 0x6d, 0x12, 0x17, 0x0a, 0x0d, 's', 'i', 'g', 'n', 'a', 't', 'u', 'r', 'e', '_', 'd', 'e', 'f',
                                0x12, 0x06, 0x0a, 0x01, 'u', 0x12, 0x01, 0x7f};
  auto md = dec.decode_metadata(mresp);
  CHECK(md && md->model_spec && md->model_spec->name == "m" && md->metadata.size() == 1);
  CHECK(md && md->metadata.size() == 1 && md->metadata[0].first == "signature_def" && md->metadata[0].second.type_url == "u" &&
        md->metadata[0].second.value == std::vector<uint8_t>{0x7f});

  std::printf(failures ? "FAILED: %d checks\n" : "cpp codec ok\n", failures);
  return failures ? 1 : 0;
}
