// tc_scene_file.hpp — a scene description on disk (flat named arrays), so C++ hosts and tests can load the
// same fixtures the Python harness uses (tests/helpers.py: save_scene_file).  In a trajopt build the
// tc_scene_desc is filled from tesseract's Environment / JointGroup instead.
//
// layout: "TCSCENE1" | u32 n_arrays | n x { u32 name_len | name | u32 dtype (0 f64, 1 i32, 2 u8) | u64 count | data }
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/trajopt_cuda.h"

namespace trajopt_cuda
{
class SceneFile
{
public:
  explicit SceneFile(const std::string& path)
  {
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f)
      throw std::runtime_error("SceneFile: cannot open " + path);
    char magic[8];
    uint32_t n = 0;
    bool ok = std::fread(magic, 1, 8, f) == 8 && std::memcmp(magic, "TCSCENE1", 8) == 0 && std::fread(&n, 4, 1, f) == 1;
    for (uint32_t i = 0; ok && i < n; ++i)
    {
      uint32_t len = 0, dtype = 0;
      uint64_t count = 0;
      ok = std::fread(&len, 4, 1, f) == 1 && len < 256;
      std::string name(len, '\0');
      ok = ok && std::fread(&name[0], 1, len, f) == len && std::fread(&dtype, 4, 1, f) == 1 &&
           std::fread(&count, 8, 1, f) == 1 && dtype <= 2;
      if (!ok)
        break;
      if (dtype == 0)
        ok = readInto(f64_[name], count, f);
      else if (dtype == 1)
        ok = readInto(i32_[name], count, f);
      else
        ok = readInto(u8_[name], count, f);
    }
    std::fclose(f);
    if (!ok)
      throw std::runtime_error("SceneFile: malformed file " + path);
  }

  // the arrays stay owned by this object: keep it alive until tc_scene_create returned
  tc_scene_desc desc(double margin, double coeff, double margin_buffer) const
  {
    tc_scene_desc d{};
    d.n_dof = scalar("n_dof");
    d.n_frames = static_cast<int32_t>(i32("frame_parent").size());
    d.frame_parent = i32("frame_parent").data();
    d.frame_joint_type = i32("frame_joint_type").data();
    d.frame_dof = i32("frame_dof").data();
    d.frame_axis = f64("frame_axis").data();
    d.frame_origin = f64("frame_origin").data();
    std::memcpy(d.base_pose, f64("base_pose").data(), sizeof(d.base_pose));
    d.n_links = static_cast<int32_t>(i32("link_frame").size());
    d.n_active_links = scalar("n_active_links");
    d.link_frame = i32("link_frame").data();
    d.link_pose = f64("link_pose").data();
    d.n_shapes = static_cast<int32_t>(i32("shape_type").size());
    d.shape_type = i32("shape_type").data();
    d.shape_link = i32("shape_link").data();
    d.shape_id = i32("shape_id").data();
    d.shape_params = f64("shape_params").data();
    d.shape_pose = f64("shape_pose").data();
    d.shape_vert_offset = i32("shape_vert_offset").data();
    d.shape_vert_count = i32("shape_vert_count").data();
    d.n_verts = static_cast<int32_t>(f64("verts").size() / 3);
    d.verts = f64("verts").data();
    d.acm = u8("acm").data();
    d.pair_margin = nullptr;
    d.pair_coeff = nullptr;
    d.default_margin = margin;
    d.default_coeff = coeff;
    d.margin_buffer = margin_buffer;
    return d;
  }
  const std::vector<double>& f64(const std::string& k) const { return get(f64_, k); }
  const std::vector<int32_t>& i32(const std::string& k) const { return get(i32_, k); }
  const std::vector<uint8_t>& u8(const std::string& k) const { return get(u8_, k); }
  bool has(const std::string& k) const { return f64_.count(k) || i32_.count(k) || u8_.count(k); }

private:
  template <class T>
  static bool readInto(std::vector<T>& v, uint64_t count, FILE* f)
  {
    v.resize(count);
    return count == 0 || std::fread(v.data(), sizeof(T), count, f) == count;
  }
  template <class T>
  static const std::vector<T>& get(const std::map<std::string, std::vector<T>>& m, const std::string& k)
  {
    auto it = m.find(k);
    if (it == m.end())
      throw std::runtime_error("SceneFile: missing array " + k);
    return it->second;
  }
  int32_t scalar(const std::string& k) const { return i32(k).at(0); }
  std::map<std::string, std::vector<double>> f64_;
  std::map<std::string, std::vector<int32_t>> i32_;
  std::map<std::string, std::vector<uint8_t>> u8_;
};
}  // namespace trajopt_cuda
