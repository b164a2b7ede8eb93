// snapshot.cpp -- see snapshot.hpp.
#include "snapshot.hpp"

#include <cstdlib>

#include "reconciler.hpp"

namespace k8sgpu_host {

// ------------------------------------------------------------------ Doc

const Doc *Doc::get(const std::string &key) const {
    if (kind != Map) return nullptr;
    for (const auto &kv : map)
        if (kv.first == key) return &kv.second;
    return nullptr;
}
std::string Doc::str(const std::string &key, const std::string &dflt) const {
    const Doc *d = get(key);
    return (d && d->kind == Scalar) ? d->scalar : dflt;
}
long Doc::num(const std::string &key, long dflt) const {
    const Doc *d = get(key);
    if (!d || d->kind != Scalar || d->scalar.empty()) return dflt;
    char *end = nullptr;
    long v = std::strtol(d->scalar.c_str(), &end, 10);
    return (end && *end == 0) ? v : dflt;
}

// ------------------------------------------------------------------ YAML subset

namespace {

struct Line {
    int indent;
    std::string text;  // without indentation, comments and trailing blanks
    int no;
};

std::string Unquote(const std::string &s) {
    if (s.size() >= 2 && ((s.front() == '"' && s.back() == '"') || (s.front() == '\'' && s.back() == '\'')))
        return s.substr(1, s.size() - 2);
    if (s == "~" || s == "null") return "";
    return s;
}

// position of the ':' that ends a key (followed by blank or end of line, outside quotes), or npos
size_t KeyColon(const std::string &t) {
    char q = 0;
    for (size_t i = 0; i < t.size(); ++i) {
        char c = t[i];
        if (q) {
            if (c == q) q = 0;
        } else if (c == '"' || c == '\'') {
            q = c;
        } else if (c == ':' && (i + 1 == t.size() || t[i + 1] == ' ')) {
            return i;
        }
    }
    return std::string::npos;
}

class YamlParser {
  public:
    YamlParser(std::vector<Line> lines, ParseError *err) : l_(std::move(lines)), err_(err) {}
    bool Block(size_t *i, int indent, Doc *out) {
        if (*i >= l_.size()) return true;
        if (l_[*i].text.compare(0, 1, "-") == 0 && (l_[*i].text.size() == 1 || l_[*i].text[1] == ' '))
            return Sequence(i, indent, out);
        return Mapping(i, indent, out);
    }

  private:
    bool Fail(const Line &ln, const std::string &m) {
        err_->msg = m;
        err_->line = ln.no;
        return false;
    }
    bool Sequence(size_t *i, int indent, Doc *out) {
        out->kind = Doc::Seq;
        while (*i < l_.size() && l_[*i].indent == indent && l_[*i].text[0] == '-' &&
               (l_[*i].text.size() == 1 || l_[*i].text[1] == ' ')) {
            std::string rest = l_[*i].text.size() > 2 ? l_[*i].text.substr(2) : "";
            size_t lead = rest.find_first_not_of(' ');
            rest = lead == std::string::npos ? "" : rest.substr(lead);
            Doc item;
            if (rest.empty()) {  // "-" alone: the item is the deeper block
                ++*i;
                if (*i < l_.size() && l_[*i].indent > indent && !Block(i, l_[*i].indent, &item)) return false;
            } else if (KeyColon(rest) != std::string::npos && rest[0] != '"' && rest[0] != '\'') {
                // "- key: value": a map whose first entry sits on the dash line
                l_[*i].indent = indent + 2 + static_cast<int>(lead == std::string::npos ? 0 : lead);
                l_[*i].text = rest;
                if (!Mapping(i, l_[*i].indent, &item)) return false;
            } else {
                item.kind = Doc::Scalar;
                item.scalar = Unquote(rest);
                ++*i;
            }
            out->seq.push_back(std::move(item));
        }
        return true;
    }
    bool Mapping(size_t *i, int indent, Doc *out) {
        out->kind = Doc::Map;
        while (*i < l_.size() && l_[*i].indent == indent) {
            const Line ln = l_[*i];
            if (ln.text[0] == '-' && (ln.text.size() == 1 || ln.text[1] == ' ')) break;  // sibling sequence of the parent
            size_t c = KeyColon(ln.text);
            if (c == std::string::npos) return Fail(ln, "expected 'key: value'");
            std::string key = Unquote(ln.text.substr(0, c));
            std::string val = c + 1 < ln.text.size() ? ln.text.substr(c + 1) : "";
            size_t lead = val.find_first_not_of(' ');
            val = lead == std::string::npos ? "" : val.substr(lead);
            Doc child;
            ++*i;
            if (!val.empty()) {
                if (val == "{}") child.kind = Doc::Map;
                else if (val == "[]") child.kind = Doc::Seq;
                else {
                    child.kind = Doc::Scalar;
                    child.scalar = Unquote(val);
                }
            } else if (*i < l_.size() && l_[*i].indent > indent) {
                if (!Block(i, l_[*i].indent, &child)) return false;
            } else if (*i < l_.size() && l_[*i].indent == indent && l_[*i].text[0] == '-' &&
                       (l_[*i].text.size() == 1 || l_[*i].text[1] == ' ')) {
                if (!Sequence(i, indent, &child)) return false;  // sequence at the key's own indent
            }
            out->map.emplace_back(std::move(key), std::move(child));
        }
        if (*i < l_.size() && l_[*i].indent > indent) return Fail(l_[*i], "unexpected indentation");
        return true;
    }
    std::vector<Line> l_;
    ParseError *err_;
};

}  // namespace

bool ParseYaml(const std::string &text, std::vector<Doc> *docs, ParseError *err) {
    std::vector<std::vector<Line>> chunks(1);
    size_t pos = 0;
    int no = 0;
    while (pos <= text.size()) {
        size_t nl = text.find('\n', pos);
        std::string raw = text.substr(pos, nl == std::string::npos ? std::string::npos : nl - pos);
        pos = nl == std::string::npos ? text.size() + 1 : nl + 1;
        ++no;
        if (!raw.empty() && raw.back() == '\r') raw.pop_back();
        // strip comments: '#' at line start or after a blank, outside quotes
        char q = 0;
        for (size_t i = 0; i < raw.size(); ++i) {
            char c = raw[i];
            if (q) {
                if (c == q) q = 0;
            } else if (c == '"' || c == '\'') {
                q = c;
            } else if (c == '#' && (i == 0 || raw[i - 1] == ' ' || raw[i - 1] == '\t')) {
                raw.erase(i);
                break;
            }
        }
        while (!raw.empty() && (raw.back() == ' ' || raw.back() == '\t')) raw.pop_back();
        if (raw.empty()) continue;
        if (raw.find('\t') != std::string::npos && raw.find_first_not_of(" \t") > raw.find('\t')) {
            err->msg = "tab in indentation";
            err->line = no;
            return false;
        }
        if (raw.compare(0, 3, "---") == 0) {
            if (!chunks.back().empty()) chunks.emplace_back();
            continue;
        }
        size_t ind = raw.find_first_not_of(' ');
        chunks.back().push_back(Line{static_cast<int>(ind), raw.substr(ind), no});
    }
    for (auto &lines : chunks) {
        if (lines.empty()) continue;
        Doc d;
        size_t i = 0;
        YamlParser p(lines, err);
        if (!p.Block(&i, lines[0].indent, &d)) return false;
        if (i != lines.size()) {
            err->msg = "unexpected dedent";
            err->line = lines[i].no;
            return false;
        }
        docs->push_back(std::move(d));
    }
    return true;
}

// ------------------------------------------------------------------ JSON

namespace {

class JsonParser {
  public:
    JsonParser(const std::string &s, ParseError *err) : s_(s), err_(err) {}
    bool Value(Doc *out) {
        Ws();
        if (p_ >= s_.size()) return Fail("unexpected end");
        char c = s_[p_];
        if (c == '{') return Object(out);
        if (c == '[') return Array(out);
        if (c == '"') {
            out->kind = Doc::Scalar;
            return String(&out->scalar);
        }
        size_t b = p_;
        while (p_ < s_.size() && std::string(",]} \n\r\t").find(s_[p_]) == std::string::npos) ++p_;
        std::string tok = s_.substr(b, p_ - b);
        if (tok.empty()) return Fail("unexpected character");
        if (tok == "null") {
            out->kind = Doc::Null;
        } else {
            out->kind = Doc::Scalar;
            out->scalar = tok;
        }
        return true;
    }
    bool AtEnd() {
        Ws();
        return p_ == s_.size();
    }

  private:
    bool Fail(const std::string &m) {
        err_->msg = m;
        err_->line = 1;
        for (size_t i = 0; i < p_ && i < s_.size(); ++i) err_->line += s_[i] == '\n';
        return false;
    }
    void Ws() {
        while (p_ < s_.size() && (s_[p_] == ' ' || s_[p_] == '\n' || s_[p_] == '\r' || s_[p_] == '\t')) ++p_;
    }
    bool String(std::string *out) {
        ++p_;  // opening quote
        while (p_ < s_.size() && s_[p_] != '"') {
            char c = s_[p_++];
            if (c == '\\') {
                if (p_ >= s_.size()) return Fail("bad escape");
                char e = s_[p_++];
                switch (e) {
                    case 'n': out->push_back('\n'); break;
                    case 't': out->push_back('\t'); break;
                    case 'r': out->push_back('\r'); break;
                    case 'b': out->push_back('\b'); break;
                    case 'f': out->push_back('\f'); break;
                    case 'u': {
                        if (p_ + 4 > s_.size()) return Fail("bad \\u escape");
                        unsigned v = static_cast<unsigned>(std::strtoul(s_.substr(p_, 4).c_str(), nullptr, 16));
                        p_ += 4;
                        if (v < 0x80) out->push_back(static_cast<char>(v));
                        else if (v < 0x800) {
                            out->push_back(static_cast<char>(0xC0 | (v >> 6)));
                            out->push_back(static_cast<char>(0x80 | (v & 0x3F)));
                        } else {
                            out->push_back(static_cast<char>(0xE0 | (v >> 12)));
                            out->push_back(static_cast<char>(0x80 | ((v >> 6) & 0x3F)));
                            out->push_back(static_cast<char>(0x80 | (v & 0x3F)));
                        }
                        break;
                    }
                    default: out->push_back(e);
                }
            } else {
                out->push_back(c);
            }
        }
        if (p_ >= s_.size()) return Fail("unterminated string");
        ++p_;
        return true;
    }
    bool Object(Doc *out) {
        out->kind = Doc::Map;
        ++p_;
        Ws();
        if (p_ < s_.size() && s_[p_] == '}') {
            ++p_;
            return true;
        }
        for (;;) {
            Ws();
            if (p_ >= s_.size() || s_[p_] != '"') return Fail("expected a string key");
            std::string key;
            if (!String(&key)) return false;
            Ws();
            if (p_ >= s_.size() || s_[p_] != ':') return Fail("expected ':'");
            ++p_;
            Doc v;
            if (!Value(&v)) return false;
            out->map.emplace_back(std::move(key), std::move(v));
            Ws();
            if (p_ < s_.size() && s_[p_] == ',') {
                ++p_;
                continue;
            }
            if (p_ < s_.size() && s_[p_] == '}') {
                ++p_;
                return true;
            }
            return Fail("expected ',' or '}'");
        }
    }
    bool Array(Doc *out) {
        out->kind = Doc::Seq;
        ++p_;
        Ws();
        if (p_ < s_.size() && s_[p_] == ']') {
            ++p_;
            return true;
        }
        for (;;) {
            Doc v;
            if (!Value(&v)) return false;
            out->seq.push_back(std::move(v));
            Ws();
            if (p_ < s_.size() && s_[p_] == ',') {
                ++p_;
                continue;
            }
            if (p_ < s_.size() && s_[p_] == ']') {
                ++p_;
                return true;
            }
            return Fail("expected ',' or ']'");
        }
    }
    const std::string &s_;
    size_t p_ = 0;
    ParseError *err_;
};

}  // namespace

bool ParseJson(const std::string &text, Doc *doc, ParseError *err) {
    JsonParser p(text, err);
    if (!p.Value(doc)) return false;
    if (!p.AtEnd()) {
        err->msg = "trailing characters after the JSON value";
        err->line = 0;
        return false;
    }
    return true;
}

// ------------------------------------------------------------------ Interner

uint32_t Interner::Id(const std::string &s) {
    auto it = ids_.find(s);
    if (it != ids_.end()) return it->second;
    uint32_t id = static_cast<uint32_t>(strs_.size());
    ids_[s] = id;
    strs_.push_back(s);
    return id;
}
bool Interner::Find(const std::string &s, uint32_t *id) const {
    auto it = ids_.find(s);
    if (it == ids_.end()) return false;
    *id = it->second;
    return true;
}

// ------------------------------------------------------------------ SnapshotBuilder

SnapshotBuilder::SnapshotBuilder() {
    RegisterSku("Standard_NC4as_T4_v3", 1, 4000, 28672);         // README.md:100,276 (4 vCPU, 28 GiB, 1x T4)
    RegisterSku("Standard_ND96isr_H100_v5", 8, 96000, 1945600);  // 96 vCPU, 1900 GiB, 8x H100
}

void SnapshotBuilder::RegisterSku(const std::string &vm_size, uint32_t cap_gpu, uint32_t cap_mcpu, uint32_t cap_mib) {
    auto it = skus_.find(vm_size);
    uint32_t id = it != skus_.end() ? it->second.sku_id : static_cast<uint32_t>(skus_.size());
    skus_[vm_size] = SkuShape{id, cap_gpu, cap_mcpu, cap_mib};
}

bool SnapshotBuilder::AddPoolsYaml(const std::string &yaml, ParseError *err) {
    std::vector<Doc> docs;
    if (!ParseYaml(yaml, &docs, err)) return false;
    for (const Doc &d : docs) {
        if (d.str("kind") != "AzureVmPool") continue;
        const Doc *meta = d.get("metadata"), *spec = d.get("spec");
        if (!meta || !spec || meta->str("name").empty()) {
            err->msg = "AzureVmPool without metadata.name or spec";
            err->line = 0;
            return false;
        }
        computev1alpha1::AzureVmPool p;
        p.Metadata.Name = meta->str("name");
        p.Metadata.Namespace = meta->str("namespace", "default");
        long rep = spec->num("replicas", -1);
        if (rep < 0 || rep > 0x7FFFFFFF) {  // +kubebuilder:validation:Minimum=0 (README.md:94)
            err->msg = "spec.replicas missing or negative in " + p.Metadata.Name;
            err->line = 0;
            return false;
        }
        p.Spec.Replicas = static_cast<int32_t>(rep);
        p.Spec.ResourceGroupName = spec->str("resourceGroupName");
        p.Spec.Location = spec->str("location");
        p.Spec.VMSize = spec->str("vmSize");
        p.Spec.VnetName = spec->str("vnetName");
        p.Spec.SubnetName = spec->str("subnetName");
        p.Spec.AzureCredentialSecret = spec->str("azureCredentialSecret");
        if (const Doc *img = spec->get("imageReference")) {
            p.Spec.ImageReference.Publisher = img->str("publisher");
            p.Spec.ImageReference.Offer = img->str("offer");
            p.Spec.ImageReference.SKU = img->str("sku");
            p.Spec.ImageReference.Version = img->str("version");
        }
        pools_.push_back(std::move(p));
    }
    return true;
}

bool SnapshotBuilder::AddVmsJson(const std::string &json, ParseError *err) {
    Doc d;
    if (!ParseJson(json, &d, err)) return false;
    if (d.kind != Doc::Seq) {
        err->msg = "VM list must be a JSON array";
        err->line = 0;
        return false;
    }
    for (const Doc &v : d.seq) {
        if (v.kind != Doc::Map || v.str("name").empty()) {
            err->msg = "VM entry without a name";
            err->line = 0;
            return false;
        }
        VM vm;
        vm.Name = v.str("name");
        if (const Doc *tags = v.get("tags"))
            for (const auto &kv : tags->map)
                if (kv.second.kind == Doc::Scalar) vm.Tags[kv.first] = kv.second.scalar;
        VmExtra ex;
        ex.location = v.str("location");
        if (const Doc *hw = v.get("hardwareProfile")) ex.vm_size = hw->str("vmSize");
        if (const Doc *z = v.get("zones"))
            if (z->kind == Doc::Seq && !z->seq.empty() && z->seq[0].kind == Doc::Scalar) ex.zone = z->seq[0].scalar;
        vms_.emplace_back(v.str("resourceGroup"), std::move(vm));
        extra_.push_back(std::move(ex));
    }
    return true;
}

bool SnapshotBuilder::ParseCpu(const std::string &s, uint32_t *mcpu) {
    if (s.empty()) return false;
    char *end = nullptr;
    if (s.back() == 'm') {
        unsigned long v = std::strtoul(s.c_str(), &end, 10);
        if (end != s.c_str() + s.size() - 1 || v >= 0x80000000ul) return false;
        *mcpu = static_cast<uint32_t>(v);
        return true;
    }
    double v = std::strtod(s.c_str(), &end);
    if (!end || *end != 0 || v < 0 || v * 1000.0 >= 2147483648.0) return false;
    *mcpu = static_cast<uint32_t>(v * 1000.0 + 0.5);
    return true;
}

bool SnapshotBuilder::ParseMemory(const std::string &s, uint32_t *mib) {
    char *end = nullptr;
    double v = std::strtod(s.c_str(), &end);
    if (!end || end == s.c_str() || v < 0) return false;
    std::string unit(end);
    double bytes;
    if (unit == "Gi") bytes = v * 1073741824.0;
    else if (unit == "Mi") bytes = v * 1048576.0;
    else if (unit == "Ki") bytes = v * 1024.0;
    else if (unit == "Ti") bytes = v * 1099511627776.0;
    else if (unit == "G") bytes = v * 1e9;
    else if (unit == "M") bytes = v * 1e6;
    else if (unit == "K" || unit == "k") bytes = v * 1e3;
    else if (unit.empty()) bytes = v;
    else return false;
    double m = (bytes + 1048575.0) / 1048576.0;  // round up: a pod must fit what it asked for
    if (m >= 2147483648.0) return false;
    *mib = static_cast<uint32_t>(m);
    return true;
}

// "gpu-1x-16c-32g-1gpu" (GPU调度平台搭建.md:534): <n>x nodes, <c> cores, <g> GiB, <k> GPUs per instance
bool SnapshotBuilder::ParseInstanceType(const std::string &s, uint32_t *gpus, uint32_t *mcpu, uint32_t *mib) {
    unsigned n = 0, c = 0, g = 0, k = 0;
    char tail = 0;
    if (std::sscanf(s.c_str(), "gpu-%ux-%uc-%ug-%ugpu%c", &n, &c, &g, &k, &tail) != 4) return false;
    if (c >= 2000000u || g >= 2000000u || k > 0x7FFFu) return false;
    *gpus = k;
    *mcpu = c * 1000u;
    *mib = g * 1024u;
    return true;
}

bool SnapshotBuilder::AddJobsYaml(const std::string &yaml, ParseError *err) {
    std::vector<Doc> docs;
    if (!ParseYaml(yaml, &docs, err)) return false;
    auto bad = [&](const std::string &m) {
        err->msg = m;
        err->line = 0;
        return false;
    };
    for (const Doc &d : docs) {
        if (const Doc *it = d.get("singleInstanceType")) {  // GoHai job template
            PodSpec p;
            p.name = d.str("name", "trainjob");
            if (it->kind != Doc::Scalar || !ParseInstanceType(it->scalar, &p.gpu, &p.mcpu, &p.mib))
                return bad("unparsable singleInstanceType");
            pods_.push_back(p);
            continue;
        }
        if (d.str("kind") != "Job") continue;
        const Doc *spec = d.get("spec");
        const Doc *tasks = spec ? spec->get("tasks") : nullptr;
        if (!tasks || tasks->kind != Doc::Seq) return bad("Job without spec.tasks");
        std::string job = d.get("metadata") ? d.get("metadata")->str("name", "job") : "job";
        for (const Doc &t : tasks->seq) {
            long replicas = t.num("replicas", 1);
            const Doc *tmpl = t.get("template");
            const Doc *pspec = tmpl ? tmpl->get("spec") : nullptr;
            const Doc *cs = pspec ? pspec->get("containers") : nullptr;
            if (!cs || cs->kind != Doc::Seq || cs->seq.empty()) return bad("task without containers");
            PodSpec p;
            for (const Doc &c : cs->seq) {  // a pod asks for the sum of its containers' limits
                const Doc *res = c.get("resources");
                const Doc *lim = res ? res->get("limits") : nullptr;
                if (!lim) continue;
                p.gpu += static_cast<uint32_t>(lim->num("nvidia.com/gpu", 0));  // GPU调度平台搭建.md:667
                uint32_t v = 0;
                if (lim->get("cpu")) {
                    if (!ParseCpu(lim->str("cpu"), &v)) return bad("unparsable cpu limit");
                    p.mcpu += v;
                }
                if (lim->get("memory")) {
                    if (!ParseMemory(lim->str("memory"), &v)) return bad("unparsable memory limit");
                    p.mib += v;
                }
            }
            if (const Doc *sel = pspec->get("nodeSelector")) {
                p.zone = sel->str("topology.kubernetes.io/zone");
                p.instance_type = sel->str("node.kubernetes.io/instance-type");
            }
            if (const Doc *tol = pspec->get("tolerations"))
                for (const Doc &x : tol->seq) {
                    std::string key = x.str("key");
                    unsigned bit = 0;
                    if (std::sscanf(key.c_str(), "gpu-taint-%u", &bit) == 1 && bit < 32) p.tol |= 1u << bit;
                }
            for (long r = 0; r < replicas; ++r) {
                p.name = job + "-" + t.str("name", "task") + "-" + std::to_string(r);
                pods_.push_back(p);
            }
        }
    }
    return true;
}

ReconcileTables SnapshotBuilder::BuildReconcileTables() const {
    ReconcileTables t;
    std::map<std::pair<std::string, std::string>, uint32_t> row;  // (resource group, owner tag) -> pool row
    for (const auto &p : pools_) {
        auto key = std::make_pair(p.Spec.ResourceGroupName, AzureVmPoolReconciler::OwnerTag(p));
        auto dup = row.find(key);
        if (dup != row.end()) {  // same owner tag as an earlier pool: its VMs cannot be told apart
            t.conflicts.emplace_back(p.Metadata.Namespace + "/" + p.Metadata.Name, t.pool_key[dup->second]);
            continue;
        }
        row[key] = static_cast<uint32_t>(t.desired.size());
        t.pool_owner.push_back(static_cast<uint32_t>(t.desired.size()));
        t.desired.push_back(p.Spec.Replicas);
        t.pool_key.push_back(p.Metadata.Namespace + "/" + p.Metadata.Name);
    }
    const uint32_t P = static_cast<uint32_t>(t.desired.size());
    t.n_owner_ids = P;
    std::vector<const computev1alpha1::AzureVmPool *> by_row(P, nullptr);
    for (const auto &p : pools_) {
        auto it = row.find(std::make_pair(p.Spec.ResourceGroupName, AzureVmPoolReconciler::OwnerTag(p)));
        if (it != row.end() && !by_row[it->second]) by_row[it->second] = &p;
    }
    for (const auto &rv : vms_) {
        const VM &vm = rv.second;
        uint32_t oid = K8SGPU_NONE_U32, ord = K8SGPU_NONE_U32;
        auto ow = vm.Tags.find(kOwnerKey);
        if (ow != vm.Tags.end()) {
            auto it = row.find(std::make_pair(rv.first, ow->second));
            if (it != row.end()) {
                oid = it->second;
                if (auto k = AzureVmPoolReconciler::OrdinalOf(vm.Name, by_row[oid]->Metadata.Name)) ord = *k;
            } else {
                oid = P;  // an owner tag no pool of this resource group carries
            }
        }
        auto mb = vm.Tags.find(kManagedByKey);
        t.vm_owner.push_back(oid);
        t.vm_flags.push_back(mb != vm.Tags.end() && mb->second == kManagedByValue ? 1u : 0u);
        t.vm_ordinal.push_back(ord);
        t.vm_name.push_back(vm.Name);
    }
    return t;
}

NodeTable SnapshotBuilder::BuildNodeTable() {
    NodeTable n;
    for (size_t i = 0; i < vms_.size(); ++i) {
        const VmExtra &ex = extra_[i];
        auto it = skus_.find(ex.vm_size);
        const bool known = it != skus_.end();
        SkuShape s = known ? it->second : SkuShape{0, 0, 0, 0};
        uint32_t zone = zones_.Id(ex.location + "-" + ex.zone);  // the topology.kubernetes.io/zone label on AKS
        n.free_gpu.push_back(s.cap_gpu);  // a freshly listed VM: nothing placed on it yet
        n.free_mcpu.push_back(s.cap_mcpu);
        n.free_mib.push_back(s.cap_mib);
        n.cap_gpu.push_back(s.cap_gpu);
        n.sku_id.push_back(s.sku_id);
        n.zone_id.push_back(zone < 32 ? zone : 31);
        n.taint_mask.push_back(0);
        n.flags.push_back(known && zone < 32 ? 1u : 0u);  // unknown size / 33rd zone: not schedulable
        n.name.push_back(vms_[i].second.Name);
    }
    return n;
}

PodTable SnapshotBuilder::BuildPodTable() {
    PodTable t;
    for (const PodSpec &p : pods_) {
        uint32_t sku_mask = K8SGPU_NONE_U32, zone_mask = K8SGPU_NONE_U32;
        if (!p.instance_type.empty()) {
            auto it = skus_.find(p.instance_type);
            sku_mask = it != skus_.end() && it->second.sku_id < 32 ? 1u << it->second.sku_id : 0u;
        }
        if (!p.zone.empty()) {
            uint32_t z = 0;
            zone_mask = zones_.Find(p.zone, &z) && z < 32 ? 1u << z : 0u;  // a zone no node is in: never feasible
        }
        t.req_gpu.push_back(p.gpu);
        t.req_mcpu.push_back(p.mcpu);
        t.req_mib.push_back(p.mib);
        t.sku_mask.push_back(sku_mask);
        t.zone_mask.push_back(zone_mask);
        t.tol_mask.push_back(p.tol);
        t.pref_zone.push_back(K8SGPU_NONE_U32);
        t.priority.push_back(0);
        t.name.push_back(p.name);
    }
    return t;
}

}  // namespace k8sgpu_host
