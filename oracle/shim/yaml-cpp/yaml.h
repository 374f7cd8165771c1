// oracle/shim/yaml-cpp/yaml.h — TEST INFRASTRUCTURE.  The reference's include/ConfigParserYAML.h only
// needs the type YAML::Node to exist (a member and five private method signatures); the parser itself
// (src/ConfigParserYAML.cpp) is not compiled — oracle/shim/ref_capi.cpp fills the tables instead.
#pragma once
namespace YAML {
class Node {};
} // namespace YAML
