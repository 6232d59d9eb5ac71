#include "nodefile.hpp"
#include "log.hpp"
#include <cctype>
#include <climits>
#include <fstream>
#include <sstream>
#include <unistd.h>

namespace ocm {

bool parse_size(const char *s, uint64_t *out) {
    if (!s || !*s) return false;
    char *end = nullptr;
    double v = strtod(s, &end);
    if (end == s || v < 0) return false;
    uint64_t mul = 1;
    switch (toupper((unsigned char)*end)) {
    case 'K': mul = 1ull << 10; ++end; break;
    case 'M': mul = 1ull << 20; ++end; break;
    case 'G': mul = 1ull << 30; ++end; break;
    case 'T': mul = 1ull << 40; ++end; break;
    default: break;
    }
    if (*end == 'i' || *end == 'I') ++end;
    if (*end == 'b' || *end == 'B') ++end;
    if (*end) return false;
    *out = (uint64_t)(v * (double)mul);
    return true;
}

bool parse_nodefile_text(const std::string &text, Nodefile *out, std::string *err) {
    std::istringstream in(text);
    std::string line;
    std::vector<NodeEntry> tmp;
    int lineno = 0;
    while (std::getline(in, line)) {
        ++lineno;
        size_t p = line.find_first_not_of(" \t\r");
        if (p == std::string::npos || line[p] == '#') continue;
        std::istringstream ls(line);
        std::vector<std::string> tok;
        for (std::string t; ls >> t;) {
            if (t[0] == '#') break;
            tok.push_back(t);
        }
        if (tok.size() < 4) {
            if (err) *err = "line " + std::to_string(lineno) + ": need `rank dns ip ocm_port [rdmacm_port] [k=v..]`";
            return false;
        }
        NodeEntry e;
        char *end = nullptr;
        long r = strtol(tok[0].c_str(), &end, 10);
        if (*end || r < 0 || r >= 4096) {
            if (err) *err = "line " + std::to_string(lineno) + ": bad rank `" + tok[0] + "`";
            return false;
        }
        e.rank = (int)r;
        e.dns = tok[1];
        e.ip_eth = tok[2];
        long port = strtol(tok[3].c_str(), &end, 10);
        if (*end || port <= 0 || port > 65535) {
            if (err) *err = "line " + std::to_string(lineno) + ": bad ocm_port `" + tok[3] + "`";
            return false;
        }
        e.ocm_port = (int)port;
        for (size_t i = 4; i < tok.size(); ++i) {
            size_t eq = tok[i].find('=');
            if (eq == std::string::npos) {
                // positional 5th column: rdmacm_port.  The reference ships 67890
                // here (> 65535); accept any integer, never use it.
                e.rdmacm_port = (int)strtol(tok[i].c_str(), nullptr, 10);
                continue;
            }
            std::string k = tok[i].substr(0, eq), v = tok[i].substr(eq + 1);
            bool ok = true;
            if (k == "gpu") e.gpu = (int)strtol(v.c_str(), nullptr, 10);
            else if (k == "gpu_bytes") ok = parse_size(v.c_str(), &e.gpu_bytes);
            else if (k == "host_bytes") ok = parse_size(v.c_str(), &e.host_bytes);
            else if (k == "numa") {
                char *ve = nullptr;
                long nn = strtol(v.c_str(), &ve, 10);
                ok = ve != v.c_str() && !*ve && nn >= -1 && nn < 1024;
                if (ok) e.numa = (int)nn;
            }
            else ok = false;
            if (!ok) {
                if (err) *err = "line " + std::to_string(lineno) + ": bad option `" + tok[i] + "`";
                return false;
            }
        }
        tmp.push_back(e);
    }
    if (tmp.empty()) {
        if (err) *err = "no node entries";
        return false;
    }
    out->nodes.assign(tmp.size(), NodeEntry());
    for (auto &e : tmp) {
        if (e.rank >= (int)tmp.size()) {
            if (err) *err = "rank " + std::to_string(e.rank) + " exceeds entry count " + std::to_string(tmp.size());
            return false;
        }
        if (out->nodes[e.rank].rank != -1) {
            if (err) *err = "duplicate rank " + std::to_string(e.rank);
            return false;
        }
        out->nodes[e.rank] = e;
    }
    return true;
}

int resolve_rank(const Nodefile &nf, const char *hostname, const char *env_rank) {
    if (env_rank && *env_rank) {
        char *end = nullptr;
        long r = strtol(env_rank, &end, 10);
        if (!*end && r >= 0 && r < nf.size()) return (int)r;
        return -1;
    }
    if (hostname)
        for (int r = nf.size() - 1; r >= 0; --r)   // highest rank first, as the reference
            if (nf.nodes[r].dns == hostname) return r;
    return nf.size() == 1 ? 0 : -1;
}

bool parse_nodefile(const std::string &path, Nodefile *out, std::string *err, int rank_override) {
    std::ifstream f(path);
    if (!f) {
        if (err) *err = "cannot open nodefile `" + path + "`";
        return false;
    }
    std::stringstream ss;
    ss << f.rdbuf();
    if (!parse_nodefile_text(ss.str(), out, err)) return false;
    if (rank_override >= 0) {
        if (rank_override >= out->size()) {
            if (err) *err = "rank override out of range";
            return false;
        }
        out->my_rank = rank_override;
        return true;
    }
    char host[HOST_NAME_MAX + 1] = {0};
    gethostname(host, HOST_NAME_MAX);
    out->my_rank = resolve_rank(*out, host, getenv("OCM_RANK"));
    if (out->my_rank < 0) {
        if (err) *err = std::string("cannot determine rank: hostname `") + host + "` not in nodefile and OCM_RANK unset/invalid";
        return false;
    }
    return true;
}

} // namespace ocm
