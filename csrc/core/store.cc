#include "core/store.h"

#include <arpa/inet.h>
#include <fcntl.h>
#include <netdb.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <poll.h>
#include <sys/socket.h>
#include <unistd.h>

#include <cerrno>
#include <cstring>
#include <stdexcept>

namespace mfl::core {

// ---------------------------------------------------------------------------------------------
// in-memory store
// ---------------------------------------------------------------------------------------------
void HashMapModelStore::InsertModel(const std::string& id, ModelBytes model) {
  std::lock_guard<std::mutex> l(mu_);
  auto& q = models_[id];
  if (policy_.lineage_length_eviction)
    while (q.size() >= policy_.Effective()) q.pop_front();
  q.push_back(std::move(model));
}

std::map<std::string, std::vector<ModelBytes>> HashMapModelStore::SelectModels(
    const std::vector<std::pair<std::string, int>>& request) {
  std::lock_guard<std::mutex> l(mu_);
  std::map<std::string, std::vector<ModelBytes>> out;
  for (auto& [id, n_req] : request) {
    auto it = models_.find(id);
    if (it == models_.end()) continue;
    const auto& q = it->second;
    auto& dst = out[id];
    if (n_req > (int)q.size()) continue;  // over-ask -> empty list
    const size_t n = n_req <= 0 ? q.size() : (size_t)n_req;
    dst.assign(q.end() - n, q.end());     // oldest -> newest
  }
  return out;
}

void HashMapModelStore::EraseModels(const std::vector<std::string>& ids) {
  std::lock_guard<std::mutex> l(mu_);
  for (auto& id : ids) models_.erase(id);
}

void HashMapModelStore::Expunge() {
  std::lock_guard<std::mutex> l(mu_);
  models_.clear();
}

int HashMapModelStore::GetLearnerLineageLength(const std::string& id) {
  std::lock_guard<std::mutex> l(mu_);
  auto it = models_.find(id);
  return it == models_.end() ? 0 : (int)it->second.size();
}

// ---------------------------------------------------------------------------------------------
// RESP client
// ---------------------------------------------------------------------------------------------
RespClient::~RespClient() { Close(); }

void RespClient::Close() {
  if (fd_ >= 0) ::close(fd_);
  fd_ = -1;
  rbuf_.clear();
  rpos_ = 0;
}

bool RespClient::Connect(const std::string& host, int port, double timeout_s, std::string* err) {
  Close();
  addrinfo hints{}, *res = nullptr;
  hints.ai_family = AF_UNSPEC;
  hints.ai_socktype = SOCK_STREAM;
  const std::string ps = std::to_string(port);
  if (getaddrinfo(host.c_str(), ps.c_str(), &hints, &res) != 0 || !res) {
    if (err) *err = "cannot resolve " + host;
    return false;
  }
  for (addrinfo* ai = res; ai; ai = ai->ai_next) {
    int fd = ::socket(ai->ai_family, ai->ai_socktype, ai->ai_protocol);
    if (fd < 0) continue;
    const int fl = fcntl(fd, F_GETFL, 0);
    fcntl(fd, F_SETFL, fl | O_NONBLOCK);
    int rc = ::connect(fd, ai->ai_addr, ai->ai_addrlen);
    if (rc != 0 && errno == EINPROGRESS) {
      pollfd p{fd, POLLOUT, 0};
      if (poll(&p, 1, (int)(timeout_s * 1000)) == 1) {
        int soerr = 0;
        socklen_t sl = sizeof(soerr);
        getsockopt(fd, SOL_SOCKET, SO_ERROR, &soerr, &sl);
        rc = soerr == 0 ? 0 : -1;
      } else {
        rc = -1;
      }
    }
    if (rc == 0) {
      fcntl(fd, F_SETFL, fl);
      int one = 1;
      setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
      timeval tv{60, 0};  // a dead server must not block the controller forever
      setsockopt(fd, SOL_SOCKET, SO_RCVTIMEO, &tv, sizeof(tv));
      setsockopt(fd, SOL_SOCKET, SO_SNDTIMEO, &tv, sizeof(tv));
      fd_ = fd;
      break;
    }
    ::close(fd);
  }
  freeaddrinfo(res);
  if (fd_ < 0 && err) *err = "cannot connect to " + host + ":" + ps;
  return fd_ >= 0;
}

bool RespClient::SendAll(const std::string& buf) {
  size_t off = 0;
  while (off < buf.size()) {
    const ssize_t n = ::send(fd_, buf.data() + off, buf.size() - off, MSG_NOSIGNAL);
    if (n <= 0) {
      if (n < 0 && errno == EINTR) continue;
      return false;
    }
    off += (size_t)n;
  }
  return true;
}

bool RespClient::ReadExact(size_t n, std::string* out) {
  while (rbuf_.size() - rpos_ < n) {
    char tmp[1 << 16];
    const ssize_t got = ::recv(fd_, tmp, sizeof(tmp), 0);
    if (got <= 0) {
      if (got < 0 && errno == EINTR) continue;
      return false;
    }
    if (rpos_ > (1u << 20) && rpos_ * 2 > rbuf_.size()) {
      rbuf_.erase(0, rpos_);
      rpos_ = 0;
    }
    rbuf_.append(tmp, (size_t)got);
  }
  out->assign(rbuf_, rpos_, n);
  rpos_ += n;
  return true;
}

bool RespClient::ReadLine(std::string* line) {
  for (;;) {
    const size_t e = rbuf_.find("\r\n", rpos_);
    if (e != std::string::npos) {
      line->assign(rbuf_, rpos_, e - rpos_);
      rpos_ = e + 2;
      return true;
    }
    char tmp[4096];
    const ssize_t got = ::recv(fd_, tmp, sizeof(tmp), 0);
    if (got <= 0) {
      if (got < 0 && errno == EINTR) continue;
      return false;
    }
    rbuf_.append(tmp, (size_t)got);
  }
}

bool RespClient::ReadReply(Reply* r) {
  std::string line;
  if (!ReadLine(&line) || line.empty()) return false;
  r->type = line[0];
  const std::string body = line.substr(1);
  switch (r->type) {
    case '+': case '-': r->str = body; return true;
    case ':': r->integer = std::stoll(body); return true;
    case '$': {
      const long long n = std::stoll(body);
      if (n < 0) { r->type = 0; return true; }
      std::string crlf;
      return ReadExact((size_t)n, &r->str) && ReadExact(2, &crlf);
    }
    case '*': {
      const long long n = std::stoll(body);
      if (n < 0) { r->type = 0; return true; }
      r->elements.resize((size_t)n);
      for (auto& e : r->elements)
        if (!ReadReply(&e)) return false;
      return true;
    }
  }
  return false;
}

bool RespClient::Command(const std::vector<std::string_view>& args, Reply* reply, std::string* err) {
  if (fd_ < 0) {
    if (err) *err = "not connected";
    return false;
  }
  std::string buf = "*" + std::to_string(args.size()) + "\r\n";
  for (auto a : args) {
    buf += "$" + std::to_string(a.size()) + "\r\n";
    buf.append(a.data(), a.size());
    buf += "\r\n";
  }
  Reply local;
  Reply* r = reply ? reply : &local;
  *r = Reply();
  if (!SendAll(buf) || !ReadReply(r)) {
    if (err) *err = "redis connection lost";
    Close();
    return false;
  }
  if (r->type == '-') {
    if (err) *err = r->str;
    return false;
  }
  return true;
}

// ---------------------------------------------------------------------------------------------
// Redis store
// ---------------------------------------------------------------------------------------------
RedisModelStore::RedisModelStore(const std::string& host, int port, EvictionPolicy p) : ModelStore(p) {
  std::string err;
  if (!client_.Connect(host, port, 1.5, &err))
    throw std::runtime_error("RedisModelStore: " + err);  // the reference exit(1)s here
}

std::string RedisModelStore::NextKey(const std::string& id) { return id + "_model_" + std::to_string(++counter_); }

void RedisModelStore::InsertModel(const std::string& id, ModelBytes model) {
  std::lock_guard<std::mutex> l(mu_);
  auto& keys = lineage_[id];
  std::string err;
  if (policy_.lineage_length_eviction) {
    while (keys.size() >= policy_.Effective()) {
      client_.Command({"DEL", keys.front()}, nullptr, &err);
      cache_.erase(keys.front());
      keys.pop_front();
    }
  }
  const std::string key = NextKey(id);
  // one LIST element per serialized Model.Variable (the framed `variables` entries, tag included)
  wire::Reader r(*model);
  std::vector<std::string_view> args{"RPUSH", key};
  while (!r.done()) {
    const uint8_t* start = r.p;
    uint32_t f; int wt; std::string_view body;
    if (!r.tag(f, wt)) break;
    if (wt == wire::kLen) {
      if (!r.bytes(body)) break;
    } else if (!r.skip(wt)) break;
    args.emplace_back(reinterpret_cast<const char*>(start), (size_t)(r.p - start));
  }
  if (args.size() == 2) args.emplace_back("");  // empty model: keep the key alive
  if (!client_.Command(args, nullptr, &err)) throw std::runtime_error("RedisModelStore RPUSH: " + err);
  keys.push_back(key);
}

std::map<std::string, std::vector<ModelBytes>> RedisModelStore::SelectModels(
    const std::vector<std::pair<std::string, int>>& request) {
  std::lock_guard<std::mutex> l(mu_);
  std::map<std::string, std::vector<ModelBytes>> out;
  std::vector<std::string> fetch;
  std::vector<std::pair<std::string, std::vector<std::string>>> plan;
  for (auto& [id, n_req] : request) {
    auto it = lineage_.find(id);
    if (it == lineage_.end()) continue;
    out[id];
    const auto& keys = it->second;
    if (n_req > (int)keys.size()) continue;
    const size_t n = n_req <= 0 ? keys.size() : (size_t)n_req;
    std::vector<std::string> sel(keys.end() - n, keys.end());
    for (auto& k : sel)
      if (!cache_.count(k)) fetch.push_back(k);
    plan.emplace_back(id, std::move(sel));
  }
  if (!fetch.empty()) {
    std::string err;
    RespClient::Reply rep;
    if (!client_.Command({"MULTI"}, nullptr, &err)) throw std::runtime_error("RedisModelStore MULTI: " + err);
    for (auto& k : fetch)
      if (!client_.Command({"LRANGE", k, "0", "-1"}, nullptr, &err))
        throw std::runtime_error("RedisModelStore LRANGE: " + err);
    if (!client_.Command({"EXEC"}, &rep, &err) || rep.elements.size() != fetch.size())
      throw std::runtime_error("RedisModelStore EXEC: " + err);
    for (size_t i = 0; i < fetch.size(); ++i) {
      auto m = std::make_shared<std::string>();
      for (auto& e : rep.elements[i].elements) m->append(e.str);
      cache_[fetch[i]] = m;
    }
  }
  for (auto& [id, sel] : plan)
    for (auto& k : sel) out[id].push_back(cache_[k]);
  return out;
}

void RedisModelStore::EraseModels(const std::vector<std::string>& ids) {
  std::lock_guard<std::mutex> l(mu_);
  std::string err;
  for (auto& id : ids) {
    auto it = lineage_.find(id);
    if (it == lineage_.end()) continue;
    for (auto& k : it->second) {
      client_.Command({"DEL", k}, nullptr, &err);
      cache_.erase(k);
    }
    lineage_.erase(it);
  }
}

void RedisModelStore::Expunge() {
  std::lock_guard<std::mutex> l(mu_);
  std::string err;
  client_.Command({"FLUSHDB"}, nullptr, &err);
  lineage_.clear();
  cache_.clear();
}

void RedisModelStore::ResetState() {
  std::lock_guard<std::mutex> l(mu_);
  cache_.clear();
}

int RedisModelStore::GetLearnerLineageLength(const std::string& id) {
  std::lock_guard<std::mutex> l(mu_);
  auto it = lineage_.find(id);
  return it == lineage_.end() ? 0 : (int)it->second.size();
}

void RedisModelStore::Shutdown() {
  std::lock_guard<std::mutex> l(mu_);
  client_.Close();
}

}  // namespace mfl::core
