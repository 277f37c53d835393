// rs_wire.hpp -- the query worker's messages on the wire, without libprotobuf.
//
// The reference's worker receives a serialised `Request` on its ZMQ SUB socket and pushes serialised
// `Reply` messages to the sink (/root/reference/src/service/readserver.proto; service.cpp:279-315,
// 1520-1577).  A batching worker has to produce those bytes for thousands of queries per kernel launch,
// so this header encodes them directly: proto2 wire format, fields in field-number order, exactly the
// bytes `Reply::SerializeToArray` writes (tests/test_wire.py compares them with google.protobuf built
// from the same schema).  No dependency beyond the standard library.
#ifndef RS_WIRE_HPP
#define RS_WIRE_HPP

#include <cstdint>
#include <string>
#include <vector>

namespace rsb200 {
namespace wire {

// ---- enums of readserver.proto -----------------------------------------------------------------
enum RequestType { CountReads = 1, ExactMatch = 2, KmerMatch = 3, SiteMatch = 4 };  // Request.t, Reply.rt
enum ReturnType { Count = 1, Reads = 2, All = 3, Samples = 4 };                     // Request.rt
enum ReplyType { ReplyCount = 1, ReplyReads = 2, ReplyAll = 3, ReplySamples = 4 };  // Reply.t

struct Request {  // readserver.proto: message Request
  int32_t t = 0, rt = 0;
  std::string q;
  bool has_k = false, has_s = false, has_p = false, has_a = false, has_isalt = false;
  int32_t k = 0, s = 0, p = 0, isalt = 0;
  std::string a;
};

struct ReadInfo {  // message ReadInfo: sample name, count, lane
  std::string g;
  int32_t c = 0, l = 0;
};

struct ResultAll {  // message ResultAll: a read and the samples that carry it
  std::string r;
  std::vector<ReadInfo> s;
};

// ---- primitives ----------------------------------------------------------------------------------
inline void put_varint(std::string& out, uint64_t v) {
  while (v >= 0x80) {
    out.push_back((char)(uint8_t)(v | 0x80));
    v >>= 7;
  }
  out.push_back((char)(uint8_t)v);
}
// int32 / enum fields: negative values are sign-extended to 64 bits (10 bytes), as protobuf does
inline void put_int32(std::string& out, uint32_t field, int32_t v) {
  put_varint(out, (uint64_t)field << 3);
  put_varint(out, (uint64_t)(int64_t)v);
}
inline void put_bytes(std::string& out, uint32_t field, const std::string& s) {
  put_varint(out, ((uint64_t)field << 3) | 2);
  put_varint(out, s.size());
  out.append(s);
}

inline bool get_varint(const uint8_t*& p, const uint8_t* end, uint64_t& v) {
  v = 0;
  for (int shift = 0; shift < 70 && p < end; shift += 7) {
    const uint8_t b = *p++;
    if (shift < 64) v |= (uint64_t)(b & 0x7F) << shift;
    if (!(b & 0x80)) return true;
  }
  return false;
}

// ---- Request -------------------------------------------------------------------------------------
inline std::string encode_request(const Request& r) {
  std::string out;
  put_int32(out, 1, r.t);
  put_int32(out, 2, r.rt);
  put_bytes(out, 3, r.q);
  if (r.has_k) put_int32(out, 4, r.k);
  if (r.has_s) put_int32(out, 5, r.s);
  if (r.has_p) put_int32(out, 6, r.p);
  if (r.has_a) put_bytes(out, 7, r.a);
  if (r.has_isalt) put_int32(out, 8, r.isalt);
  return out;
}

// Request::ParseFromArray: unknown fields are skipped; false on malformed input or a missing required field.
inline bool decode_request(const void* data, size_t size, Request& r) {
  const uint8_t* p = (const uint8_t*)data;
  const uint8_t* end = p + size;
  r = Request();
  bool has_t = false, has_rt = false, has_q = false;
  while (p < end) {
    uint64_t key, v;
    if (!get_varint(p, end, key)) return false;
    const uint32_t field = (uint32_t)(key >> 3), wt = (uint32_t)(key & 7);
    if (wt == 0) {
      if (!get_varint(p, end, v)) return false;
      const int32_t x = (int32_t)(uint32_t)v;
      switch (field) {
        case 1: r.t = x; has_t = true; break;
        case 2: r.rt = x; has_rt = true; break;
        case 4: r.k = x; r.has_k = true; break;
        case 5: r.s = x; r.has_s = true; break;
        case 6: r.p = x; r.has_p = true; break;
        case 8: r.isalt = x; r.has_isalt = true; break;
        default: break;
      }
    } else if (wt == 2) {
      if (!get_varint(p, end, v) || v > (uint64_t)(end - p)) return false;
      if (field == 3) { r.q.assign((const char*)p, (size_t)v); has_q = true; }
      else if (field == 7) { r.a.assign((const char*)p, (size_t)v); r.has_a = true; }
      p += v;
    } else if (wt == 1) {
      if (end - p < 8) return false;
      p += 8;
    } else if (wt == 5) {
      if (end - p < 4) return false;
      p += 4;
    } else {
      return false;
    }
  }
  return has_t && has_rt && has_q;
}

// ---- Reply ---------------------------------------------------------------------------------------
inline std::string encode_read_info(const ReadInfo& i) {
  std::string m;
  put_bytes(m, 1, i.g);
  put_int32(m, 2, i.c);
  put_int32(m, 3, i.l);
  return m;
}

inline void reply_head(std::string& out, int32_t rt, int32_t t, const std::string& q) {
  put_int32(out, 1, rt);
  put_int32(out, 2, t);
  put_bytes(out, 3, q);
}

// count_reads (service.cpp:279-315): one message per strand, c { forward_matches | revcomp_matches { c } };
// ResultCount.c is int32: the 64-bit interval width is truncated exactly as set_c() does.
inline std::string encode_reply_count(int32_t rt, const std::string& q, bool revcomp, uint64_t count) {
  std::string rc_msg;  // ResultCount
  put_int32(rc_msg, 1, (int32_t)(uint32_t)count);
  std::string c_msg;  // ReplyCount
  put_bytes(c_msg, revcomp ? 2 : 1, rc_msg);
  std::string out;
  reply_head(out, rt, ReplyCount, q);
  put_bytes(out, 4, c_msg);
  return out;
}

// output=reads (service.cpp:714-797): r { repeated forward_matches | revcomp_matches { r } }
inline std::string encode_reply_reads(int32_t rt, const std::string& q, bool revcomp, const std::vector<std::string>& reads) {
  std::string r_msg;  // ReplyReads
  for (const std::string& rd : reads) {
    std::string one;  // ResultReads
    put_bytes(one, 1, rd);
    put_bytes(r_msg, revcomp ? 2 : 1, one);
  }
  std::string out;
  reply_head(out, rt, ReplyReads, q);
  put_bytes(out, 5, r_msg);
  return out;
}

// output=all (service.cpp:799-860): a { repeated forward_matches | revcomp_matches { r, repeated s } }
// QueryTask sets Reply.t to the request's return type and fills `a` for everything but Reads (service.cpp:1263,
// 1292): a request with rt = Samples is answered with t = 4 and field `a`; hence reply_t.
inline std::string encode_reply_all(int32_t rt, const std::string& q, bool revcomp, const std::vector<ResultAll>& hits,
                                    int32_t reply_t = ReplyAll) {
  std::string a_msg;  // ReplyAll
  for (const ResultAll& h : hits) {
    std::string one;  // ResultAll
    put_bytes(one, 1, h.r);
    for (const ReadInfo& i : h.s) put_bytes(one, 2, encode_read_info(i));
    put_bytes(a_msg, revcomp ? 2 : 1, one);
  }
  std::string out;
  reply_head(out, rt, reply_t, q);
  put_bytes(out, 6, a_msg);
  return out;
}

// output=samples (service_samples.cpp:85-118): s { repeated s }
inline std::string encode_reply_samples(int32_t rt, const std::string& q, const std::vector<ReadInfo>& samples) {
  std::string s_msg;  // ResultSamples
  for (const ReadInfo& i : samples) put_bytes(s_msg, 1, encode_read_info(i));
  std::string out;
  reply_head(out, rt, ReplySamples, q);
  put_bytes(out, 7, s_msg);
  return out;
}

}  // namespace wire
}  // namespace rsb200

#endif
