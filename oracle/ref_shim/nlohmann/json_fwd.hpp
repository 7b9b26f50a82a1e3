#include <nlohmann/json.hpp>
