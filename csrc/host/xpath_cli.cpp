// Command-line edits of the case file:  XPATH = v | XPATH @attr = v | XPATH inject|insert [first|last|before|after] <xml/>
// | XPATH delete | XPATH print | -s out.xml | -x   (reference src/xpath_modification.cpp:4-161)
#include "../core/common.h"
#include "xml.h"

namespace tclb {

int xpath_modify(XmlDocument &doc, XmlNode *config, const std::vector<std::string> &a) {
    const int n = (int)a.size();
    for (int i = 0; i < n; i++) {
        if (a[i] == "-s") {
            if (++i >= n) { logf(L_FATAL, "no filename for -s"); return -1; }
            doc.save_file(a[i]);
            continue;
        }
        if (a[i] == "-x") { logf(L_OUTPUT, "Gracefully exiting"); return -444; }
        logf(L_OUTPUT, "XPATH: %s", a[i].c_str());
        std::vector<XPathResult> found;
        try { found = xpath_select(config, a[i]); }
        catch (const Error &e) {
            logf(L_FATAL, "XPATH: parsing error: %s", e.what());
            logf(L_FATAL, "XPATH: Syntax: .../tclb model file.xml XPATH = value (spaces are important)");
            return -1;
        }
        logf(L_OUTPUT, "XPATH: %zu things found", found.size());
        if (++i >= n) { logf(L_FATAL, "no operator in xpath evaluation"); return -1; }
        std::string add_attr;
        if (a[i][0] == '@') { add_attr = a[i].substr(1); if (++i >= n) { logf(L_FATAL, "no operator in xpath evaluation"); return -1; } }
        const std::string &op = a[i];
        if (op == "=") {
            if (++i >= n) { logf(L_FATAL, "XPATH: No value supplied to = operator"); return -1; }
            if (found.empty()) { logf(L_FATAL, "XPATH: Nothing selected for substitution"); return -1; }
            for (auto &r : found) {
                if (add_attr.empty()) {
                    if (r.attr.empty()) {
                        logf(L_FATAL, "XPATH: Operator = can only be used for attributes");
                        logf(L_FATAL, " eg. Model/Param/@value = 0 - change attribute");
                        logf(L_FATAL, "     Model/Param @value = 0 - add or change attribute");
                        return -1;
                    }
                    r.node->set_attr(r.attr, a[i]);
                    logf(L_OUTPUT, "XPATH: Set attr %s to \"%s\" in <%s />", r.attr.c_str(), a[i].c_str(), r.node->name.c_str());
                } else {
                    if (!r.attr.empty()) { logf(L_FATAL, "XPATH: '@attr =' needs a node selection"); return -1; }
                    bool had = r.node->has_attr(add_attr);
                    r.node->set_attr(add_attr, a[i]);
                    logf(L_OUTPUT, "XPATH: %s attr %s = \"%s\" in <%s />", had ? "Set" : "Adding", add_attr.c_str(), a[i].c_str(), r.node->name.c_str());
                }
            }
        } else if (op == "inject" || op == "insert") {
            int type = 0;
            if (++i >= n) { logf(L_FATAL, "XPATH: No value or specifier supplied to inject operator"); return -1; }
            if (a[i] == "last") { type = 0; i++; } else if (a[i] == "first") { type = 1; i++; }
            else if (a[i] == "after") { type = 2; i++; } else if (a[i] == "before") { type = 3; i++; }
            if (i >= n) { logf(L_FATAL, "XPATH: No value supplied to inject operator"); return -1; }
            std::unique_ptr<XmlNode> frag;
            try {
                XmlDocument d;
                d.parse(a[i]);
                if (d.children().size() != 1) { logf(L_FATAL, "XPATH: exactly one XML element expected in inject string"); return -1; }
                frag = d.detach(d.root());
            } catch (const Error &e) { logf(L_FATAL, "XPATH: Error while parsing inject string: %s", e.what()); return -1; }
            if (found.empty()) { logf(L_FATAL, "XPATH: Nothing selected for injection"); return -1; }
            if (found.size() != 1) logf(L_WARNING_IMP, "XPATH: More then one thing selected for injection");
            XPathResult &r = found[0];
            if (!r.attr.empty()) { logf(L_FATAL, "XPATH: Operator 'insert' can only be used for nodes"); return -1; }
            if (type == 0) r.node->append(std::move(frag));
            else if (type == 1) r.node->prepend(std::move(frag));
            else if (type == 2) r.node->parent->insert_after(std::move(frag), r.node);
            else r.node->parent->insert_before(std::move(frag), r.node);
        } else if (op == "delete") {
            for (auto &r : found) {
                if (r.attr.empty()) { logf(L_OUTPUT, "XPATH: Removed node: <%s />", r.node->name.c_str()); if (r.node->parent) r.node->parent->remove(r.node); }
                else { logf(L_OUTPUT, "XPATH: Removed attribute: %s", r.attr.c_str()); r.node->remove_attr(r.attr); }
            }
        } else if (op == "print") {
            for (auto &r : found) {
                if (r.attr.empty()) logf(L_OUTPUT, "XPATH: Node: %s", r.node->name.c_str());
                else logf(L_OUTPUT, "XPATH: Attr: %s=\"%s\"", r.attr.c_str(), r.node->attr(r.attr).c_str());
            }
        } else {
            logf(L_FATAL, "Unknown operator in xpath evaluation: %s", op.c_str());
            logf(L_FATAL, "Currently supported: =, print, insert/inject, delete");
            return -1;
        }
    }
    return 0;
}
}  // namespace tclb
