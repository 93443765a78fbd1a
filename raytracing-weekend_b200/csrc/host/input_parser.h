// input_parser.h — token-vector command line lookup.
// Same interface as the reference's InputParser (InputParser.h:9-30):
// getCmdOption(opt) returns the token after `opt` or "", cmdOptionExists(opt).
#pragma once
#include <algorithm>
#include <string>
#include <vector>

class InputParser {
public:
    InputParser(int argc, char** argv) : tokens_(argv + (argc > 0 ? 1 : 0), argv + argc) {}
    const std::string& getCmdOption(const std::string& option) const {
        static const std::string empty;
        auto it = std::find(tokens_.begin(), tokens_.end(), option);
        if (it != tokens_.end() && ++it != tokens_.end()) return *it;
        return empty;
    }
    bool cmdOptionExists(const std::string& option) const {
        return std::find(tokens_.begin(), tokens_.end(), option) != tokens_.end();
    }

private:
    std::vector<std::string> tokens_;
};
