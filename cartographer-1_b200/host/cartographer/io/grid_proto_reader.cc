#include "cartographer/io/grid_proto_reader.h"

#include <limits>
#include <vector>

#include "cartographer/io/proto_wire.h"

namespace cartographer {
namespace io {
namespace {

using wire::Reader;
using wire::Span;

struct ParsedLimits {
  double resolution = 0., max_x = 0., max_y = 0.;
  int num_x_cells = 0, num_y_cells = 0;
};

// transform.proto Vector2d {double x = 1; double y = 2;}
void ParseVector2d(const Span& s, double* x, double* y) {
  Reader r(s);
  int field, type;
  while (r.Next(&field, &type)) {
    if (field == 1 && type == 1) *x = r.Double();
    else if (field == 2 && type == 1) *y = r.Double();
    else r.Skip(type);
  }
  CSM_CHECK(r.ok());
}
// cell_limits_2d.proto CellLimits {int32 num_x_cells = 1; int32 num_y_cells = 2;}
void ParseCellLimits(const Span& s, ParsedLimits* out) {
  Reader r(s);
  int field, type;
  while (r.Next(&field, &type)) {
    if (field == 1 && type == 0) out->num_x_cells = static_cast<int32_t>(r.Varint());
    else if (field == 2 && type == 0) out->num_y_cells = static_cast<int32_t>(r.Varint());
    else r.Skip(type);
  }
  CSM_CHECK(r.ok());
}
// map_limits.proto MapLimits {double resolution = 1; Vector2d max = 2; CellLimits cell_limits = 3;}
ParsedLimits ParseMapLimits(const Span& s) {
  ParsedLimits out;
  Reader r(s);
  int field, type;
  while (r.Next(&field, &type)) {
    if (field == 1 && type == 1) out.resolution = r.Double();
    else if (field == 2 && type == 2) ParseVector2d(r.Bytes(), &out.max_x, &out.max_y);
    else if (field == 3 && type == 2) ParseCellLimits(r.Bytes(), &out);
    else r.Skip(type);
  }
  CSM_CHECK(r.ok());
  return out;
}

// The reference's CHECK_LE(cell, std::numeric_limits<uint16>::max()) on an int32 element.
uint16 CheckedCell(const uint64_t varint) {
  const int32_t cell = static_cast<int32_t>(varint);
  CSM_CHECK(cell >= 0 && cell <= std::numeric_limits<uint16>::max());
  return static_cast<uint16>(cell);
}

}  // namespace

std::unique_ptr<mapping::Grid2D> ReadGrid2DProto(const void* data, size_t size,
                                                 mapping::ValueConversionTables* const conversion_tables) {
  // grid_2d.proto: MapLimits limits = 1; repeated int32 cells = 2; CellBox known_cells_box = 3;
  // oneof grid {ProbabilityGrid probability_grid_2d = 4; TSDF2D tsdf_2d = 5;}
  // float min_correspondence_cost = 6; float max_correspondence_cost = 7;
  ParsedLimits limits;
  bool has_limits = false, has_probability_grid = false, has_tsdf = false;
  std::vector<uint16> cells, weight_cells;
  float min_cost = 0.f, max_cost = 0.f, truncation_distance = 0.f, max_weight = 0.f;
  Reader r(data, size);
  int field, type;
  while (r.Next(&field, &type)) {
    if (field == 1 && type == 2) {
      limits = ParseMapLimits(r.Bytes());
      has_limits = true;
    } else if (field == 2) {
      wire::RepeatedVarint(&r, type, [&cells](uint64_t v) { cells.push_back(CheckedCell(v)); });
    } else if (field == 4 && type == 2) {
      r.Bytes();  // message ProbabilityGrid {}
      has_probability_grid = true;
      has_tsdf = false;  // oneof: the last one on the wire wins
    } else if (field == 5 && type == 2) {
      // tsdf_2d.proto: float truncation_distance = 1; float max_weight = 2; repeated int32 weight_cells = 3;
      Reader t(r.Bytes());
      weight_cells.clear();
      int f, ty;
      while (t.Next(&f, &ty)) {
        if (f == 1 && ty == 5) truncation_distance = t.Float();
        else if (f == 2 && ty == 5) max_weight = t.Float();
        else if (f == 3) wire::RepeatedVarint(&t, ty, [&weight_cells](uint64_t v) { weight_cells.push_back(CheckedCell(v)); });
        else t.Skip(ty);
      }
      CSM_CHECK(t.ok());
      has_tsdf = true;
      has_probability_grid = false;
    } else if (field == 6 && type == 5) {
      min_cost = r.Float();
    } else if (field == 7 && type == 5) {
      max_cost = r.Float();
    } else {
      r.Skip(type);  // known_cells_box (3) only speeds up cropping; unknown fields are skipped
    }
  }
  CSM_CHECK(r.ok());
  CSM_CHECK(has_limits);
  CSM_CHECK(has_probability_grid || has_tsdf);  // ProbabilityGrid / TSDF2D constructors CHECK has_*()
  // grid_2d.cc:22-44: both zero means an older file format, load the default bounds.
  if (min_cost == 0.f && max_cost == 0.f) {
    min_cost = mapping::kMinCorrespondenceCost;
    max_cost = mapping::kMaxCorrespondenceCost;
  }
  CSM_CHECK(min_cost < max_cost);
  const mapping::MapLimits map_limits(limits.resolution, transform::Vector2d(limits.max_x, limits.max_y),
                                      mapping::CellLimits(limits.num_x_cells, limits.num_y_cells));
  const size_t n = static_cast<size_t>(limits.num_x_cells) * limits.num_y_cells;
  CSM_CHECK(cells.size() == n);  // the reference indexes cells by ToFlatIndex without a check
  std::unique_ptr<mapping::Grid2D> grid;
  if (has_probability_grid) {
    grid.reset(new mapping::ProbabilityGrid(map_limits, min_cost, max_cost, conversion_tables));
  } else {
    CSM_CHECK(weight_cells.size() == n);
    auto* tsdf = new mapping::TSDF2D(map_limits, min_cost, max_cost, truncation_distance, max_weight, conversion_tables);
    *tsdf->mutable_weight_cells() = std::move(weight_cells);
    grid.reset(tsdf);
  }
  *grid->mutable_correspondence_cost_cells() = std::move(cells);
  return grid;
}

std::unique_ptr<mapping::HybridGrid> ReadHybridGridProto(const void* data, size_t size) {
  // hybrid_grid.proto: float resolution = 1; repeated sint32 x_indices = 3, y_indices = 4,
  // z_indices = 5; repeated int32 values = 6;
  float resolution = 0.f;
  std::vector<int32_t> x, y, z, values;
  Reader r(data, size);
  int field, type;
  while (r.Next(&field, &type)) {
    if (field == 1 && type == 5) resolution = r.Float();
    else if (field == 3) wire::RepeatedVarint(&r, type, [&x](uint64_t v) { x.push_back(Reader::ZigZag32(v)); });
    else if (field == 4) wire::RepeatedVarint(&r, type, [&y](uint64_t v) { y.push_back(Reader::ZigZag32(v)); });
    else if (field == 5) wire::RepeatedVarint(&r, type, [&z](uint64_t v) { z.push_back(Reader::ZigZag32(v)); });
    else if (field == 6) wire::RepeatedVarint(&r, type, [&values](uint64_t v) { values.push_back(static_cast<int32_t>(v)); });
    else r.Skip(type);
  }
  CSM_CHECK(r.ok());
  CSM_CHECK(resolution > 0.f);
  CSM_CHECK(values.size() == x.size() && values.size() == y.size() && values.size() == z.size());
  std::unique_ptr<mapping::HybridGrid> grid(new mapping::HybridGrid(resolution));
  for (size_t i = 0; i != values.size(); ++i) {
    // hybrid_grid.h:479-483: SetProbability(index, ValueToProbability(value)); ValueToProbability reads
    // a 32768-entry table (an out-of-range value is undefined behaviour there, an abort here) and
    // value 0 (unknown) becomes kMinProbability, i.e. value 1.
    CSM_CHECK(values[i] >= 0 && values[i] < 32768);
    const float scale = (mapping::kMaxProbability - mapping::kMinProbability) / 32766.f;
    const float probability = values[i] == 0 ? mapping::kMinProbability
                                             : values[i] * scale + (mapping::kMinProbability - scale);
    grid->SetProbability(Eigen::Array3i(x[i], y[i], z[i]), probability);
  }
  return grid;
}

}  // namespace io
}  // namespace cartographer
