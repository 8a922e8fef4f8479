// Minimal XML DOM with the XPath subset the case-file tooling needs.
// The reference vendors pugixml (src/pugixml.cpp, 14.7 kLoC third-party); this is a compact
// purpose-built replacement: parser (elements, attributes, text, comments, CDATA, PIs), serializer,
// tree editing, and XPath location paths with / // * . .. @attr comment() text() node() and
// predicates [n] [last()] [@a] [@a='v'] [name] [name='v'] joined by `and`/`or`.
#pragma once
#include <memory>
#include <string>
#include <utility>
#include <vector>

namespace tclb {

enum class XmlType { Document, Element, Text, Comment, CData, PI };

class XmlNode {
  public:
    XmlType type;
    std::string name;       // element / PI name
    std::string value;      // text, comment, cdata, PI content
    std::vector<std::pair<std::string, std::string>> attrs;
    std::vector<std::unique_ptr<XmlNode>> kids;
    XmlNode *parent = nullptr;

    explicit XmlNode(XmlType t = XmlType::Element, const std::string &n = "") : type(t), name(n) {}
    // attributes
    bool has_attr(const std::string &a) const;
    std::string attr(const std::string &a, const std::string &def = "") const;
    void set_attr(const std::string &a, const std::string &v);
    bool remove_attr(const std::string &a);
    bool attr_bool(const std::string &a, bool def = false) const;
    // children
    XmlNode *child(const std::string &n) const;
    std::vector<XmlNode *> children(const std::string &n = "") const;     // elements only
    XmlNode *next_sibling_element() const;
    XmlNode *append(std::unique_ptr<XmlNode> k);
    XmlNode *append_element(const std::string &n);
    XmlNode *prepend(std::unique_ptr<XmlNode> k);
    XmlNode *insert_before(std::unique_ptr<XmlNode> k, const XmlNode *ref);
    XmlNode *insert_after(std::unique_ptr<XmlNode> k, const XmlNode *ref);
    std::unique_ptr<XmlNode> detach(XmlNode *k);
    bool remove(XmlNode *k);
    std::unique_ptr<XmlNode> clone() const;
    std::string text() const;                                            // concatenated text children
    std::string path() const;                                            // /a/b[2]/c
    // serialisation
    std::string to_string(int indent = 0, bool pretty = true) const;
};

class XmlDocument : public XmlNode {
  public:
    XmlDocument() : XmlNode(XmlType::Document) {}
    void parse(const std::string &text);              // throws tclb::Error with line number
    void load_file(const std::string &file);
    void save_file(const std::string &file) const;
    XmlNode *root() const;                            // first element
    std::string to_string_doc() const;
};
std::unique_ptr<XmlNode> xml_parse_fragment(const std::string &text);   // single element

struct XPathResult {
    XmlNode *node = nullptr;
    std::string attr;                                 // non-empty when the step selected an attribute
};
std::vector<XPathResult> xpath_select(XmlNode *context, const std::string &expr);

std::string xml_escape(const std::string &s, bool attr);

}  // namespace tclb
