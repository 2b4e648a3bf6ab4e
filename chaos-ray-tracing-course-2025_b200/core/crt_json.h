#pragma once
// Scene loader, same entry point as the reference (src/core/crt_json.h:11).
#include <filesystem>
#include <istream>
#include <optional>
#include <string_view>

#include "crt_scene.h"

namespace crt::json {

// nullopt on any validation failure, no exceptions, no message -- exactly the reference's contract.
std::optional<Scene> read_scene_from_istream(std::istream &is, const std::filesystem::path &asset_root);
// same loader on an in-memory document (what the C ABI and _crt use; avoids the stream copy)
std::optional<Scene> read_scene_from_memory(std::string_view text, const std::filesystem::path &asset_root);

} // namespace crt::json
