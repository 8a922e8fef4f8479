#include "xml.h"
#include "../core/common.h"
#include <algorithm>
#include <cctype>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <sstream>

namespace tclb {

// ---------------------------------------------------------------- node API ----
bool XmlNode::has_attr(const std::string &a) const {
    for (auto &p : attrs) if (p.first == a) return true;
    return false;
}
std::string XmlNode::attr(const std::string &a, const std::string &def) const {
    for (auto &p : attrs) if (p.first == a) return p.second;
    return def;
}
void XmlNode::set_attr(const std::string &a, const std::string &v) {
    for (auto &p : attrs) if (p.first == a) { p.second = v; return; }
    attrs.emplace_back(a, v);
}
bool XmlNode::remove_attr(const std::string &a) {
    for (size_t i = 0; i < attrs.size(); i++) if (attrs[i].first == a) { attrs.erase(attrs.begin() + i); return true; }
    return false;
}
bool XmlNode::attr_bool(const std::string &a, bool def) const {
    if (!has_attr(a)) return def;
    std::string v = attr(a);
    return !v.empty() && (v[0] == '1' || v[0] == 't' || v[0] == 'T' || v[0] == 'y' || v[0] == 'Y');
}
XmlNode *XmlNode::child(const std::string &n) const {
    for (auto &k : kids) if (k->type == XmlType::Element && k->name == n) return k.get();
    return nullptr;
}
std::vector<XmlNode *> XmlNode::children(const std::string &n) const {
    std::vector<XmlNode *> out;
    for (auto &k : kids) if (k->type == XmlType::Element && (n.empty() || k->name == n)) out.push_back(k.get());
    return out;
}
XmlNode *XmlNode::next_sibling_element() const {
    if (!parent) return nullptr;
    bool seen = false;
    for (auto &k : parent->kids) {
        if (seen && k->type == XmlType::Element) return k.get();
        if (k.get() == this) seen = true;
    }
    return nullptr;
}
XmlNode *XmlNode::append(std::unique_ptr<XmlNode> k) {
    k->parent = this;
    kids.push_back(std::move(k));
    return kids.back().get();
}
XmlNode *XmlNode::append_element(const std::string &n) { return append(std::make_unique<XmlNode>(XmlType::Element, n)); }
XmlNode *XmlNode::prepend(std::unique_ptr<XmlNode> k) {
    k->parent = this;
    kids.insert(kids.begin(), std::move(k));
    return kids.front().get();
}
XmlNode *XmlNode::insert_before(std::unique_ptr<XmlNode> k, const XmlNode *ref) {
    for (size_t i = 0; i < kids.size(); i++)
        if (kids[i].get() == ref) {
            k->parent = this;
            kids.insert(kids.begin() + i, std::move(k));
            return kids[i].get();
        }
    return append(std::move(k));
}
XmlNode *XmlNode::insert_after(std::unique_ptr<XmlNode> k, const XmlNode *ref) {
    for (size_t i = 0; i < kids.size(); i++)
        if (kids[i].get() == ref) {
            k->parent = this;
            kids.insert(kids.begin() + i + 1, std::move(k));
            return kids[i + 1].get();
        }
    return append(std::move(k));
}
std::unique_ptr<XmlNode> XmlNode::detach(XmlNode *k) {
    for (size_t i = 0; i < kids.size(); i++)
        if (kids[i].get() == k) {
            std::unique_ptr<XmlNode> r = std::move(kids[i]);
            kids.erase(kids.begin() + i);
            r->parent = nullptr;
            return r;
        }
    return nullptr;
}
bool XmlNode::remove(XmlNode *k) { return detach(k) != nullptr; }
std::unique_ptr<XmlNode> XmlNode::clone() const {
    auto n = std::make_unique<XmlNode>(type, name);
    n->value = value;
    n->attrs = attrs;
    for (auto &k : kids) n->append(k->clone());
    return n;
}
std::string XmlNode::text() const {
    std::string s;
    for (auto &k : kids) if (k->type == XmlType::Text || k->type == XmlType::CData) s += k->value;
    return s;
}
std::string XmlNode::path() const {
    if (!parent) return "";
    int idx = 0, cnt = 0;
    for (auto &k : parent->kids)
        if (k->type == XmlType::Element && k->name == name) { cnt++; if (k.get() == this) idx = cnt; }
    std::string s = parent->path() + "/" + name;
    if (cnt > 1) s += "[" + std::to_string(idx) + "]";
    return s;
}
std::string xml_escape(const std::string &s, bool attr) {
    std::string o;
    for (char c : s) {
        switch (c) {
        case '&': o += "&amp;"; break;
        case '<': o += "&lt;"; break;
        case '>': o += "&gt;"; break;
        case '"': if (attr) o += "&quot;"; else o += c; break;
        default: o += c;
        }
    }
    return o;
}
std::string XmlNode::to_string(int indent, bool pretty) const {
    std::string pad = pretty ? std::string(indent, '\t') : "", nl = pretty ? "\n" : "";
    switch (type) {
    case XmlType::Text: return pad + xml_escape(value, false) + nl;
    case XmlType::CData: return pad + "<![CDATA[" + value + "]]>" + nl;
    case XmlType::Comment: return pad + "<!--" + value + "-->" + nl;
    case XmlType::PI: return pad + "<?" + name + (value.empty() ? "" : " " + value) + "?>" + nl;
    case XmlType::Document: {
        std::string s;
        for (auto &k : kids) s += k->to_string(0, pretty);
        return s;
    }
    case XmlType::Element: break;
    }
    std::string s = pad + "<" + name;
    for (auto &a : attrs) s += " " + a.first + "=\"" + xml_escape(a.second, true) + "\"";
    if (kids.empty()) return s + " />" + nl;
    bool only_text = true;
    for (auto &k : kids) if (k->type != XmlType::Text && k->type != XmlType::CData) only_text = false;
    if (only_text) {
        s += ">";
        for (auto &k : kids) s += k->type == XmlType::Text ? xml_escape(k->value, false) : "<![CDATA[" + k->value + "]]>";
        return s + "</" + name + ">" + nl;
    }
    s += ">" + nl;
    for (auto &k : kids) s += k->to_string(indent + 1, pretty);
    return s + pad + "</" + name + ">" + nl;
}

// ---------------------------------------------------------------- parser ----
namespace {
struct Parser {
    const std::string &s;
    size_t i = 0;
    explicit Parser(const std::string &t) : s(t) {}
    int line() const { return 1 + (int)std::count(s.begin(), s.begin() + std::min(i, s.size()), '\n'); }
    [[noreturn]] void err(const std::string &m) const { fail("XML parse error at line %d: %s", line(), m.c_str()); }
    bool starts(const char *p) const { return s.compare(i, strlen(p), p) == 0; }
    void skip_ws() { while (i < s.size() && isspace((unsigned char)s[i])) i++; }
    static bool name_char(char c) { return isalnum((unsigned char)c) || c == '_' || c == '-' || c == '.' || c == ':'; }
    std::string name() {
        size_t a = i;
        while (i < s.size() && name_char(s[i])) i++;
        if (i == a) err("name expected");
        return s.substr(a, i - a);
    }
    std::string unescape(const std::string &t) {
        std::string o;
        for (size_t k = 0; k < t.size(); k++) {
            if (t[k] != '&') { o += t[k]; continue; }
            size_t e = t.find(';', k);
            if (e == std::string::npos) { o += t[k]; continue; }
            std::string ent = t.substr(k + 1, e - k - 1);
            if (ent == "amp") o += '&';
            else if (ent == "lt") o += '<';
            else if (ent == "gt") o += '>';
            else if (ent == "quot") o += '"';
            else if (ent == "apos") o += '\'';
            else if (!ent.empty() && ent[0] == '#') {
                long c = ent.size() > 1 && (ent[1] == 'x' || ent[1] == 'X') ? strtol(ent.c_str() + 2, nullptr, 16) : strtol(ent.c_str() + 1, nullptr, 10);
                if (c < 0x80) o += (char)c;
                else if (c < 0x800) { o += (char)(0xC0 | (c >> 6)); o += (char)(0x80 | (c & 0x3F)); }
                else { o += (char)(0xE0 | (c >> 12)); o += (char)(0x80 | ((c >> 6) & 0x3F)); o += (char)(0x80 | (c & 0x3F)); }
            } else { o += "&" + ent + ";"; }
            k = e;
        }
        return o;
    }
    void parse_content(XmlNode *parent) {
        while (i < s.size()) {
            if (s[i] != '<') {
                size_t a = i;
                while (i < s.size() && s[i] != '<') i++;
                std::string t = s.substr(a, i - a);
                if (t.find_first_not_of(" \t\r\n") != std::string::npos) {
                    auto n = std::make_unique<XmlNode>(XmlType::Text);
                    n->value = unescape(trim(t));
                    parent->append(std::move(n));
                }
                continue;
            }
            if (starts("</")) return;
            if (starts("<!--")) {
                size_t e = s.find("-->", i + 4);
                if (e == std::string::npos) err("unterminated comment");
                auto n = std::make_unique<XmlNode>(XmlType::Comment);
                n->value = s.substr(i + 4, e - i - 4);
                parent->append(std::move(n));
                i = e + 3;
            } else if (starts("<![CDATA[")) {
                size_t e = s.find("]]>", i + 9);
                if (e == std::string::npos) err("unterminated CDATA");
                auto n = std::make_unique<XmlNode>(XmlType::CData);
                n->value = s.substr(i + 9, e - i - 9);
                parent->append(std::move(n));
                i = e + 3;
            } else if (starts("<?")) {
                size_t e = s.find("?>", i + 2);
                if (e == std::string::npos) err("unterminated processing instruction");
                std::string body = s.substr(i + 2, e - i - 2);
                auto n = std::make_unique<XmlNode>(XmlType::PI);
                size_t sp = body.find_first_of(" \t\r\n");
                n->name = body.substr(0, sp);
                n->value = sp == std::string::npos ? "" : trim(body.substr(sp));
                parent->append(std::move(n));
                i = e + 2;
            } else if (starts("<!")) {       // DOCTYPE etc.: skip (with nested brackets)
                int depth = 0;
                while (i < s.size()) {
                    if (s[i] == '[') depth++;
                    if (s[i] == ']') depth--;
                    if (s[i] == '>' && depth <= 0) { i++; break; }
                    i++;
                }
            } else parse_element(parent);
        }
    }
    void parse_element(XmlNode *parent) {
        i++;   // <
        XmlNode *n = parent->append_element(name());
        while (true) {
            skip_ws();
            if (i >= s.size()) err("unterminated start tag <" + n->name + ">");
            if (starts("/>")) { i += 2; return; }
            if (s[i] == '>') { i++; break; }
            std::string an = name();
            skip_ws();
            if (i >= s.size() || s[i] != '=') err("'=' expected after attribute " + an);
            i++;
            skip_ws();
            if (i >= s.size() || (s[i] != '"' && s[i] != '\'')) err("quoted attribute value expected for " + an);
            char q = s[i++];
            size_t e = s.find(q, i);
            if (e == std::string::npos) err("unterminated attribute value");
            n->attrs.emplace_back(an, unescape(s.substr(i, e - i)));
            i = e + 1;
        }
        parse_content(n);
        if (!starts("</")) err("missing end tag </" + n->name + ">");
        i += 2;
        std::string en = name();
        if (en != n->name) err("end tag </" + en + "> does not match <" + n->name + ">");
        skip_ws();
        if (i >= s.size() || s[i] != '>') err("'>' expected");
        i++;
    }
};
}  // namespace

void XmlDocument::parse(const std::string &text) {
    kids.clear();
    Parser p(text);
    p.parse_content(this);
    if (p.i < text.size()) p.err("unexpected end tag");
    if (!root()) fail("XML document has no root element");
}
void XmlDocument::load_file(const std::string &file) {
    std::ifstream f(file, std::ios::binary);
    if (!f) fail("cannot open %s", file.c_str());
    std::stringstream ss;
    ss << f.rdbuf();
    try { parse(ss.str()); }
    catch (const Error &e) { fail("%s: %s", file.c_str(), e.what()); }
}
std::string XmlDocument::to_string_doc() const {
    std::string s;
    bool has_decl = !kids.empty() && kids[0]->type == XmlType::PI && kids[0]->name == "xml";
    if (!has_decl) s = "<?xml version=\"1.0\"?>\n";
    return s + to_string(0, true);
}
void XmlDocument::save_file(const std::string &file) const {
    make_parent_dirs(file);
    std::ofstream f(file, std::ios::binary);
    if (!f) fail("cannot write %s", file.c_str());
    f << to_string_doc();
}
XmlNode *XmlDocument::root() const {
    for (auto &k : kids) if (k->type == XmlType::Element) return k.get();
    return nullptr;
}
std::unique_ptr<XmlNode> xml_parse_fragment(const std::string &text) {
    XmlDocument d;
    d.parse(text);
    return d.detach(d.root());
}

// ---------------------------------------------------------------- XPath subset ----
namespace {
struct Step {
    bool descendant = false;   // preceded by //
    enum Kind { Name, Any, Comment, Text, Node, Self, Parent, Attr } kind = Name;
    std::string name;
    std::vector<std::string> preds;
};
std::vector<Step> parse_path(const std::string &e, bool &absolute) {
    std::vector<Step> steps;
    size_t i = 0;
    absolute = false;
    bool desc = false;
    if (e.compare(0, 2, "//") == 0) { absolute = true; desc = true; i = 2; }
    else if (!e.empty() && e[0] == '/') { absolute = true; i = 1; }
    while (i < e.size()) {
        Step st;
        st.descendant = desc;
        desc = false;
        size_t a = i;
        int depth = 0;
        char quote = 0;
        while (i < e.size()) {
            char c = e[i];
            if (quote) { if (c == quote) quote = 0; }
            else if (c == '\'' || c == '"') quote = c;
            else if (c == '[') depth++;
            else if (c == ']') depth--;
            else if (c == '/' && depth == 0) break;
            i++;
        }
        std::string tok = e.substr(a, i - a);
        size_t b = tok.find('[');
        std::string nt = trim(b == std::string::npos ? tok : tok.substr(0, b));
        while (b != std::string::npos) {
            int d = 0;
            size_t k = b;
            char q = 0;
            for (; k < tok.size(); k++) {
                char c = tok[k];
                if (q) { if (c == q) q = 0; continue; }
                if (c == '\'' || c == '"') q = c;
                else if (c == '[') d++;
                else if (c == ']') { d--; if (d == 0) break; }
            }
            if (k >= tok.size()) fail("XPath: unbalanced [ in '%s'", e.c_str());
            st.preds.push_back(trim(tok.substr(b + 1, k - b - 1)));
            b = tok.find('[', k);
        }
        if (nt == "*") st.kind = Step::Any;
        else if (nt == ".") st.kind = Step::Self;
        else if (nt == "..") st.kind = Step::Parent;
        else if (nt == "comment()") st.kind = Step::Comment;
        else if (nt == "text()") st.kind = Step::Text;
        else if (nt == "node()") st.kind = Step::Node;
        else if (!nt.empty() && nt[0] == '@') { st.kind = Step::Attr; st.name = nt.substr(1); }
        else if (nt.compare(0, 7, "child::") == 0) st.name = nt.substr(7);
        else if (nt.compare(0, 11, "attribute::") == 0) { st.kind = Step::Attr; st.name = nt.substr(11); }
        else if (nt.empty()) fail("XPath: empty step in '%s'", e.c_str());
        else st.name = nt;
        steps.push_back(st);
        if (i < e.size()) {
            if (e.compare(i, 2, "//") == 0) { desc = true; i += 2; }
            else i += 1;
        }
    }
    return steps;
}
bool match(const XmlNode *n, const Step &st) {
    switch (st.kind) {
    case Step::Name: return n->type == XmlType::Element && n->name == st.name;
    case Step::Any: return n->type == XmlType::Element;
    case Step::Comment: return n->type == XmlType::Comment;
    case Step::Text: return n->type == XmlType::Text || n->type == XmlType::CData;
    case Step::Node: return true;
    default: return false;
    }
}
std::string unquote(const std::string &s) {
    std::string t = trim(s);
    if (t.size() >= 2 && (t[0] == '\'' || t[0] == '"') && t.back() == t[0]) return t.substr(1, t.size() - 2);
    return t;
}
bool eval_simple(const XmlNode *n, const std::string &p, int pos, int last);
bool eval_pred(const XmlNode *n, const std::string &p, int pos, int last) {
    // split on top-level " or " then " and "
    auto split_kw = [](const std::string &s, const std::string &kw) {
        std::vector<std::string> out;
        char q = 0;
        size_t a = 0;
        for (size_t i = 0; i < s.size(); i++) {
            if (q) { if (s[i] == q) q = 0; continue; }
            if (s[i] == '\'' || s[i] == '"') { q = s[i]; continue; }
            if (s.compare(i, kw.size(), kw) == 0) { out.push_back(s.substr(a, i - a)); a = i + kw.size(); i = a - 1; }
        }
        out.push_back(s.substr(a));
        return out;
    };
    for (auto &alt : split_kw(p, " or ")) {
        bool all = true;
        for (auto &c : split_kw(alt, " and ")) if (!eval_simple(n, trim(c), pos, last)) { all = false; break; }
        if (all) return true;
    }
    return false;
}
bool eval_simple(const XmlNode *n, const std::string &p, int pos, int last) {
    if (p.empty()) return true;
    if (p == "last()") return pos == last;
    if (std::all_of(p.begin(), p.end(), [](char c) { return isdigit((unsigned char)c); })) return pos == atoi(p.c_str());
    if (p.compare(0, 10, "position()") == 0) {
        size_t e = p.find('=');
        return e != std::string::npos && pos == atoi(p.c_str() + e + 1);
    }
    if (p.compare(0, 4, "not(") == 0 && p.back() == ')') return !eval_pred(n, p.substr(4, p.size() - 5), pos, last);
    std::string lhs = p, rhs;
    bool has_eq = false, neq = false;
    char q = 0;
    for (size_t i = 0; i < p.size(); i++) {
        if (q) { if (p[i] == q) q = 0; continue; }
        if (p[i] == '\'' || p[i] == '"') { q = p[i]; continue; }
        if (p[i] == '!' && i + 1 < p.size() && p[i + 1] == '=') { lhs = trim(p.substr(0, i)); rhs = unquote(p.substr(i + 2)); has_eq = neq = true; break; }
        if (p[i] == '=') { lhs = trim(p.substr(0, i)); rhs = unquote(p.substr(i + 1)); has_eq = true; break; }
    }
    std::string val;
    bool exists = false;
    if (!lhs.empty() && lhs[0] == '@') { exists = n->has_attr(lhs.substr(1)); val = n->attr(lhs.substr(1)); }
    else if (lhs == "text()" || lhs == ".") { exists = true; val = n->text(); }
    else if (lhs == "name()") { exists = true; val = n->name; }
    else { XmlNode *c = n->child(lhs); exists = c != nullptr; if (c) val = c->text(); }
    if (!has_eq) return exists;
    return exists && ((val == rhs) != neq);
}
void descend(XmlNode *n, std::vector<XmlNode *> &out) {
    out.push_back(n);
    for (auto &k : n->kids) descend(k.get(), out);
}
}  // namespace

std::vector<XPathResult> xpath_select(XmlNode *context, const std::string &expr) {
    std::vector<XPathResult> results;
    // union
    {
        std::vector<std::string> parts;
        char q = 0;
        int depth = 0;
        size_t a = 0;
        for (size_t i = 0; i < expr.size(); i++) {
            char c = expr[i];
            if (q) { if (c == q) q = 0; continue; }
            if (c == '\'' || c == '"') q = c;
            else if (c == '[') depth++;
            else if (c == ']') depth--;
            else if (c == '|' && depth == 0) { parts.push_back(expr.substr(a, i - a)); a = i + 1; }
        }
        if (!parts.empty()) {
            parts.push_back(expr.substr(a));
            for (auto &p : parts) { auto r = xpath_select(context, trim(p)); results.insert(results.end(), r.begin(), r.end()); }
            return results;
        }
    }
    bool absolute = false;
    std::vector<Step> steps = parse_path(trim(expr), absolute);
    XmlNode *start = context;
    if (absolute) while (start->parent) start = start->parent;
    std::vector<XmlNode *> cur{start};
    for (size_t si = 0; si < steps.size(); si++) {
        const Step &st = steps[si];
        if (st.kind == Step::Attr) {
            if (si + 1 != steps.size()) fail("XPath: attribute step must be last in '%s'", expr.c_str());
            std::vector<XmlNode *> base;
            if (st.descendant) for (auto *n : cur) descend(n, base); else base = cur;
            for (auto *n : base) if (n->type == XmlType::Element && (st.name == "*" || n->has_attr(st.name))) {
                if (st.name == "*") for (auto &a : n->attrs) results.push_back({n, a.first});
                else results.push_back({n, st.name});
            }
            return results;
        }
        std::vector<XmlNode *> next;
        for (auto *n : cur) {
            std::vector<XmlNode *> cand;
            if (st.kind == Step::Self) cand.push_back(n);
            else if (st.kind == Step::Parent) { if (n->parent) cand.push_back(n->parent); }
            else if (st.descendant) {
                std::vector<XmlNode *> all;
                for (auto &k : n->kids) descend(k.get(), all);
                for (auto *d : all) if (match(d, st)) cand.push_back(d);
            } else {
                for (auto &k : n->kids) if (match(k.get(), st)) cand.push_back(k.get());
            }
            if (st.descendant && !st.preds.empty()) {
                // positional predicates are relative to each parent
                std::vector<XmlNode *> filtered;
                for (auto *c : cand) {
                    int pos = 0, last = 0, my = 0;
                    if (c->parent) for (auto &k : c->parent->kids) if (match(k.get(), st)) { last++; if (k.get() == c) my = last; }
                    pos = my;
                    bool ok = true;
                    for (auto &p : st.preds) if (!eval_pred(c, p, pos, last)) { ok = false; break; }
                    if (ok) filtered.push_back(c);
                }
                cand = filtered;
            } else {
                for (auto &p : st.preds) {
                    std::vector<XmlNode *> filtered;
                    int last = (int)cand.size(), pos = 0;
                    for (auto *c : cand) { pos++; if (eval_pred(c, p, pos, last)) filtered.push_back(c); }
                    cand = filtered;
                }
            }
            for (auto *c : cand) if (std::find(next.begin(), next.end(), c) == next.end()) next.push_back(c);
        }
        cur = next;
    }
    for (auto *n : cur) results.push_back({n, ""});
    return results;
}
}  // namespace tclb
