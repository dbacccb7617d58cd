// velociraptor-stf_b200/csrc/fof.cuh -- device-resident state shared by the FOF / unbinding stages.
#pragma once

#include <functional>
#include "common.cuh"

namespace vrgpu {

// Particles as uploaded from the caller's (strided) view, original order.
struct DeviceParticles {
    size_t n = 0;
    DBuf<double> pos;      // packed xyz, 3n
    DBuf<double> vel;      // packed xyz, 3n (optional)
    DBuf<double> mass;     // n (optional)
    DBuf<long long> pid;   // n (optional)
    bool has_vel = false, has_mass = false, has_pid = false;
};

// Uniform fine-cell grid used by the 3D linker.  Cell edge c < b/sqrt(3) so that any two particles in the
// same cell are linked by construction; neighbours are searched within +-2 cells.
struct Grid {
    double lo[3];
    double inv_c[3];
    double c[3];
    int nc[3];
    double period; // 0 = open
    int periodic[3]; // per axis: wraps at `period`
    int bits[3];   // bits per axis in the packed key
};

struct Fof3dState {
    size_t n = 0;
    uint32_t ncells = 0;
    Grid grid{};
    DBuf<uint32_t> order;      // sorted slot -> original index
    DBuf<double> sx, sy, sz;   // positions in sorted (cell) order
    DBuf<uint32_t> cell_of;    // sorted slot -> cell
    DBuf<uint32_t> cell_start; // ncells + 1
    DBuf<uint64_t> cell_key;   // ncells
    DBuf<uint32_t> root;       // ncells: root cell after linking
    DBuf<long long> pfof;      // n, by ORIGINAL index
    long long ngroups = 0;
    long long ngrouped = 0;
};

// helpers in api.cu / io.cu
void upload_view(vrgpu_ctx *ctx, const vrgpu_view &view, bool need_vel, bool need_mass, bool need_pid, DeviceParticles &out);
void download_i64(vrgpu_ctx *ctx, const long long *dev, long long *host, size_t n);
// device -> host in page-locked chunks; consume(chunk, byte offset, length, part, nparts) runs on up to 8 host threads per
// chunk while the next chunk is in flight; chunk boundaries fall on multiples of `granule` bytes
void staged_download(vrgpu_ctx *ctx, const void *dev, size_t bytes, size_t granule,
                     const std::function<void(const unsigned char *, size_t, size_t, int, int)> &consume);

// stages
void fof3d_device(vrgpu_ctx *ctx, const DeviceParticles &P, double b, double period, long long minsize, Fof3dState &S);
// the two halves of fof3d_device: linking (periodic_axes == nullptr: every axis wraps when period > 0) and labelling
void fof3d_link(vrgpu_ctx *ctx, const DeviceParticles &P, double b, double period, const int *periodic_axes, Fof3dState &S);
void fof3d_label(vrgpu_ctx *ctx, const DeviceParticles &P, long long minsize, Fof3dState &S);
// AdjustStructureForPeriod on the device copy (P.pos updated in place); returns the number of moved particles and
// fills compact lists (original index, new xyz) for the host write-back.
size_t adjust_for_period_device(vrgpu_ctx *ctx, DeviceParticles &P, const long long *pfof_dev, long long ngroups, double period,
                                DBuf<uint32_t> &moved_idx, DBuf<double> &moved_xyz, bool by_pid = false);

} // namespace vrgpu
