// TEST INFRASTRUCTURE (oracle/_ref build only): stand-in for protoc's output of
// submap_visualization.proto (SubmapTexture is what Grid2D::DrawToSubmapTexture fills).
#ifndef ORACLE_REF_SHIM_SUBMAP_VISUALIZATION_PB_H_
#define ORACLE_REF_SHIM_SUBMAP_VISUALIZATION_PB_H_
#include "cartographer/transform/proto/transform.pb.h"
namespace cartographer {
namespace mapping {
namespace proto {
class SubmapQuery_Response_SubmapTexture {
  SHIM_STRING(cells)
  SHIM_SCALAR(int32_t, width)
  SHIM_SCALAR(int32_t, height)
  SHIM_SCALAR(double, resolution)
  SHIM_MESSAGE(transform::proto::Rigid3d, slice_pose)
};
class SubmapQuery_Response {
 public:
  typedef SubmapQuery_Response_SubmapTexture SubmapTexture;
  SHIM_SCALAR(int32_t, submap_version)
  SHIM_REPEATED(SubmapTexture, textures)
  SHIM_STRING(error_message)
};
class SubmapQuery {
 public:
  typedef SubmapQuery_Response Response;
};
}  // namespace proto
}  // namespace mapping
}  // namespace cartographer
#endif
