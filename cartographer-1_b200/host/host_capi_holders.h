// Handle types shared by the flat C entry points (host_capi*.cc).
#ifndef CSM_HOST_CAPI_HOLDERS_H_
#define CSM_HOST_CAPI_HOLDERS_H_

#include <memory>

#include "cartographer_compat.h"
#include "csm_b200.h"

// A 2D grid handle: the grid and the conversion tables it borrows.  A grid that belongs to a
// loaded mapping state (csmh_state_submap_grid) aliases the state's shared_ptr instead, which
// keeps the state -- and the tables its grids borrow -- alive as long as the handle.
struct GridHolder {
  cartographer::mapping::ValueConversionTables tables;
  std::shared_ptr<cartographer::mapping::Grid2D> grid;
  const void* locked = nullptr;  // the cells' storage while it is page-locked (csmh_grid_pin)
  ~GridHolder() {
    if (locked != nullptr) csm_host_unregister(const_cast<void*>(locked));
  }
};

#endif  // CSM_HOST_CAPI_HOLDERS_H_
