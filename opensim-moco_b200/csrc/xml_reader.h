// Minimal XML DOM reader for OpenSim documents (.osim, ExternalLoads .xml).
// Only what the flattener needs: elements, attributes, text, comments, the declaration.
#pragma once
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace mocob200 {

struct XmlNode {
    std::string tag;
    std::vector<std::pair<std::string, std::string>> attrs;
    std::string text;
    std::vector<std::unique_ptr<XmlNode>> children;

    const std::string* attr(const std::string& k) const {
        for (auto& a : attrs)
            if (a.first == k) return &a.second;
        return nullptr;
    }
    std::string name() const {
        auto* n = attr("name");
        return n ? *n : std::string();
    }
    const XmlNode* child(const std::string& t) const {
        for (auto& c : children)
            if (c->tag == t) return c.get();
        return nullptr;
    }
    std::vector<const XmlNode*> childrenOf() const {
        std::vector<const XmlNode*> v;
        for (auto& c : children) v.push_back(c.get());
        return v;
    }
    // text of a direct child element, trimmed; `def` if absent
    std::string childText(const std::string& t, const std::string& def = "") const;
    bool has(const std::string& t) const { return child(t) != nullptr; }
    // depth-first search for all descendants with this tag
    void findAll(const std::string& t, std::vector<const XmlNode*>& out) const {
        for (auto& c : children) {
            if (c->tag == t) out.push_back(c.get());
            c->findAll(t, out);
        }
    }
};

std::unique_ptr<XmlNode> parseXmlFile(const std::string& path);
std::unique_ptr<XmlNode> parseXmlString(const std::string& s);
std::string trim(const std::string& s);
std::vector<double> parseDoubles(const std::string& s);

} // namespace mocob200
