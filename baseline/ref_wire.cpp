// Wire-format interoperability harness over the UNMODIFIED reference's BinarySerializer / Message / TBuffer
// (tests/test_reference_interop.py).  Frames are built exactly as TcpCommunicator::send_message does
// (tcp_communicator.hpp:190-205): FixedHeader(message.size()) then the message, in one TBuffer.
//
//   ref_wire write <dir>          writes none.frame, bool.frame, string.frame, job.frame (a 2x3x4x5 tensor of i*0.25-3)
//   ref_wire read <frame file>    parses a frame produced elsewhere and prints its fields, one per line
//   ref_wire stageconfig <json>   StageConfig::from_json on a configuration produced elsewhere, prints its to_json()
#include <cstdio>
#include <fstream>
#include <iterator>
#include <string>
#include <vector>

#include "pipeline/binary_serializer.hpp"
#include "pipeline/message.hpp"
#include "pipeline/stage_config.hpp"
#include "pipeline/tbuffer.hpp"

using namespace tnn;

static void dump(const std::string &path, Message &&m, const std::string &sender) {
  m.header().sender_id = sender;
  TBuffer buf;
  FixedHeader fh(m.size());
  BinarySerializer::serialize(fh, buf);
  BinarySerializer::serialize(m, buf);
  std::ofstream f(path, std::ios::binary);
  f.write(reinterpret_cast<const char *>(buf.get()), buf.size());
}

int main(int argc, char **argv) {
  if (argc < 3) return 2;
  const std::string mode = argv[1];
  if (mode == "write") {
    const std::string d = argv[2];
    dump(d + "/none.frame", Message("stage_0", CommandType::TRAIN_MODE), "coordinator");
    dump(d + "/bool.frame", Message("coordinator", CommandType::CONFIG_RECEIVED, true), "stage_1");
    dump(d + "/string.frame", Message("stage_1", CommandType::CONFIG_TRANSFER, std::string("{\"stage_id\": \"stage_1\"}")),
         "coordinator");
    Tensor<float> t({2, 3, 4, 5});
    for (size_t i = 0; i < t.size(); ++i) t.data()[i] = 0.25f * (float)i - 3.f;
    Job<float> job;
    job.data = std::move(t);
    job.micro_batch_id = 7;
    dump(d + "/job.frame", Message("stage_1", CommandType::FORWARD_JOB, std::move(job)), "stage_0");
    return 0;
  }
  if (mode == "stageconfig") {
    std::ifstream f(argv[2]);
    nlohmann::json j;
    f >> j;
    StageConfig cfg = StageConfig::from_json(j);
    std::printf("stage_id %s\nnext %s %s\nprev %s %s\nroundtrip %s\n", cfg.stage_id.c_str(),
                cfg.next_stage_endpoint.get_parameter<std::string>("host").c_str(),
                cfg.next_stage_endpoint.get_parameter<std::string>("port").c_str(),
                cfg.prev_stage_endpoint.get_parameter<std::string>("host").c_str(),
                cfg.prev_stage_endpoint.get_parameter<std::string>("port").c_str(), cfg.to_json().dump().c_str());
    return 0;
  }
  if (mode == "read") {
    std::ifstream f(argv[2], std::ios::binary);
    std::vector<uint8_t> raw((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
    TBuffer buf(raw.size());
    buf.resize(raw.size());
    memcpy(buf.get(), raw.data(), raw.size());
    size_t off = 0;
    FixedHeader fh;
    BinarySerializer::deserialize(buf, off, fh);
    Message m;
    BinarySerializer::deserialize(buf, off, m);
    std::printf("version %d\nlength %llu\nconsumed %zu\nrecipient %s\nsender %s\ncommand %d\npayload_type %llu\n",
                (int)fh.PROTOCOL_VERSION, (unsigned long long)fh.length, off, m.header().recipient_id.c_str(),
                m.header().sender_id.c_str(), (int)m.header().command_type, (unsigned long long)m.data().payload_type);
    if (m.has_type<Job<float>>()) {
      auto &job = m.get<Job<float>>();
      std::printf("micro_batch %zu\nshape", job.micro_batch_id);
      for (size_t d : job.data.shape()) std::printf(" %zu", d);
      double s = 0;
      for (size_t i = 0; i < job.data.size(); ++i) s += (double)job.data.data()[i] * (double)(i % 7 + 1);
      std::printf("\nchecksum %.9g\n", s);
    } else if (m.has_type<std::string>()) {
      std::printf("string %s\n", m.get<std::string>().c_str());
    } else if (m.has_type<bool>()) {
      std::printf("bool %d\n", (int)m.get<bool>());
    }
    return 0;
  }
  return 2;
}
