"""readserver.proto (/root/reference/src/service/readserver.proto) rebuilt as dynamic google.protobuf classes:
protoc is not in the image, so the schema is restated field by field here.  TEST INFRASTRUCTURE: it is the
independent encoder the C++ wire codec (include/rs_wire.hpp) is compared with."""
from google.protobuf import descriptor_pb2, descriptor_pool, message_factory

_F = descriptor_pb2.FieldDescriptorProto
REQ, OPT, REP = _F.LABEL_REQUIRED, _F.LABEL_OPTIONAL, _F.LABEL_REPEATED


def _msg(fd, name, fields, enums=()):
    m = fd.message_type.add()
    m.name = name
    for ename, values in enums:
        e = m.enum_type.add()
        e.name = ename
        for vname, num in values:
            v = e.value.add()
            v.name, v.number = vname, num
    for fname, num, label, typ, tname in fields:
        f = m.field.add()
        f.name, f.number, f.label, f.type = fname, num, label, typ
        if tname:
            f.type_name = tname
    return m


def build():
    fd = descriptor_pb2.FileDescriptorProto()
    fd.name, fd.syntax = "readserver_restated.proto", "proto2"
    I32, STR, MSG, ENUM = _F.TYPE_INT32, _F.TYPE_STRING, _F.TYPE_MESSAGE, _F.TYPE_ENUM
    rtypes = [("CountReads", 1), ("ExactMatch", 2), ("KmerMatch", 3), ("SiteMatch", 4)]
    _msg(fd, "Request", [("t", 1, REQ, ENUM, ".Request.RequestType"), ("rt", 2, REQ, ENUM, ".Request.ReturnType"),
                         ("q", 3, REQ, STR, None), ("k", 4, OPT, I32, None), ("s", 5, OPT, I32, None),
                         ("p", 6, OPT, I32, None), ("a", 7, OPT, STR, None), ("isalt", 8, OPT, I32, None)],
         enums=[("RequestType", rtypes), ("ReturnType", [("Count", 1), ("Reads", 2), ("All", 3), ("Samples", 4)])])
    _msg(fd, "ReadInfo", [("g", 1, REQ, STR, None), ("c", 2, REQ, I32, None), ("l", 3, REQ, I32, None)])
    _msg(fd, "ResultSamples", [("s", 1, REP, MSG, ".ReadInfo")])
    _msg(fd, "ResultAll", [("r", 1, REQ, STR, None), ("s", 2, REP, MSG, ".ReadInfo")])
    _msg(fd, "ResultCount", [("c", 1, REQ, I32, None)])
    _msg(fd, "ResultReads", [("r", 1, REQ, STR, None)])
    _msg(fd, "Reply", [("rt", 1, REQ, ENUM, ".Reply.RequestType"), ("t", 2, REQ, ENUM, ".Reply.ReplyType"),
                       ("q", 3, REQ, STR, None), ("c", 4, OPT, MSG, ".ReplyCount"), ("r", 5, OPT, MSG, ".ReplyReads"),
                       ("a", 6, OPT, MSG, ".ReplyAll"), ("s", 7, OPT, MSG, ".ResultSamples")],
         enums=[("RequestType", rtypes),
                ("ReplyType", [("ReplyCount", 1), ("ReplyReads", 2), ("ReplyAll", 3), ("ResultSamples", 4)])])
    _msg(fd, "ReplyAll", [("forward_matches", 1, REP, MSG, ".ResultAll"), ("revcomp_matches", 2, REP, MSG, ".ResultAll")])
    _msg(fd, "ReplyCount", [("forward_matches", 1, OPT, MSG, ".ResultCount"), ("revcomp_matches", 2, OPT, MSG, ".ResultCount")])
    _msg(fd, "ReplyReads", [("forward_matches", 1, REP, MSG, ".ResultReads"), ("revcomp_matches", 2, REP, MSG, ".ResultReads")])
    pool = descriptor_pool.DescriptorPool()
    pool.Add(fd)
    get = lambda n: message_factory.GetMessageClass(pool.FindMessageTypeByName(n))
    return {n: get(n) for n in ("Request", "Reply", "ReadInfo", "ResultAll", "ResultCount", "ResultReads")}


def count_reply(M, rt: int, q: str, revcomp: bool, count: int) -> bytes:
    """What count_reads (service.cpp:279-315) serialises: ResultCount.c is int32 (set_c truncates)."""
    r = M["Reply"]()
    r.rt, r.t, r.q = rt, 1, q
    c = ((count & 0xFFFFFFFF) ^ 0x80000000) - 0x80000000
    (r.c.revcomp_matches if revcomp else r.c.forward_matches).c = c
    return r.SerializeToString()
