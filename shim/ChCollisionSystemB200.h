// ChCollisionSystemB200 -- the collision-system seam of Chrono::Multicore over libchrono_b200 (include/chrono_b200.h).
//
// In a Chrono build this class derives from ChCollisionSystemChronoMulticore and overrides the virtuals of
// chrono::ChCollisionSystem (src/chrono/collision/ChCollisionSystem.h:41-231) that
// ChSystemMulticore::AdvanceDynamics calls (src/chrono_multicore/physics/ChSystemMulticore.cpp:93-101): PreProcess, Run,
// PostProcess, GetBoundingBox, GetTimerCollisionBroad / Narrow, ResetTimers, and the lazy fill of the contact arrays that
// ReportContacts and AssembleSystem read.  The base class itself needs Eigen (absent here), so this file
// compiles the SAME method bodies against the reference's real data contract, chrono::ChCollisionData
// (src/chrono/collision/multicore/ChCollisionData.h:89-222), which is all those methods touch.  With CHRONO_B200_IN_TREE
// defined the class declaration gains the base and the `override`s; nothing else changes.
#pragma once
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "chrono/collision/multicore/ChCollisionData.h"
#include "chrono_b200.h"

namespace chrono {

#ifdef CHRONO_B200_IN_TREE
#define B200_BASE : public ChCollisionSystemChronoMulticore
#define B200_OVERRIDE override
#else
#define B200_BASE
#define B200_OVERRIDE
#endif

class ChCollisionSystemB200 B200_BASE {
  public:
    /// Owns one cb200_system on `device`; `data` is the collision data of the system (what
    /// ChCollisionSystemChronoMulticore shares with the data manager, ChCollisionSystemChronoMulticore.cpp:25-35).
    ChCollisionSystemB200(std::shared_ptr<ChCollisionData> data, int device = 0);
    ~ChCollisionSystemB200();

    // ---- settings that live in ChBroadphase / ChNarrowphase in the reference (ChCollisionSystemMulticore.h:120-160)
    void SetBroadphaseGridResolution(const vec3& bins);            // GridType::FIXED_RESOLUTION
    void SetBroadphaseGridSize(const real3& bin_size);             // GridType::FIXED_BIN_SIZE
    void SetBroadphaseGridDensity(real density);                   // GridType::FIXED_DENSITY
    void SetNarrowphaseAlgorithm(int algorithm);                   // ChNarrowphase::Algorithm (MPR 0, PRIMS 1, HYBRID 2)
    void SetEnvelope(real envelope) { cd_data->collision_envelope = envelope; }
    /// ChCollisionSystemMulticore::SetActiveBoundingBox (ChCollisionSystemMulticore.cpp:72-77); the flags the box clears
    /// on the device are written back to state_data.active_rigid in PostProcess, where the reference's Run leaves them.
    void SetActiveBoundingBox(const real3& aabbmin, const real3& aabbmax);

    // ---- the ChCollisionSystem virtuals on the hot path
    void PreProcess() B200_OVERRIDE;   ///< new shapes / bodies and the current pose go to the device
    void Run() B200_OVERRIDE;          ///< GenerateAABB + broadphase + narrowphase (ChCollisionSystemMulticore.cpp:286-322)
    void PostProcess() B200_OVERRIDE;  ///< counters and grid back into cd_data (measures, :324-345)
    void GetBoundingBox(real3& aabb_min, real3& aabb_max) const;
    double GetTimerCollisionBroad() const B200_OVERRIDE;
    double GetTimerCollisionNarrow() const B200_OVERRIDE;
    void ResetTimers() B200_OVERRIDE;

    /// The contact arrays of cd_data (norm / cpta / cptb / dpth / erad / bids _rigid_rigid, contact_shapeIDs) are filled
    /// on demand: ReportContacts, ChSystemMulticoreNSC::AssembleSystem and user code read them, a step does not.
    void FetchContacts();
    /// pair_shapeIDs (candidate pairs), bin_number / bin_aabb_number / bin_active / bin_start_index and the AABBs
    void FetchBroadphase();

    cb200_system* Handle() { return sys; }

  private:
    void PushSettings();
    void PushNewBodiesAndShapes();
    void UploadState();
    static void Check(int rc) {
        if (rc != CB200_OK)
            throw std::runtime_error(std::string("libchrono_b200: ") + cb200_last_error());
    }

    std::shared_ptr<ChCollisionData> cd_data;
    cb200_system* sys = nullptr;
    cb200_settings settings;
    bool use_aabb_active = false;
    size_t bodies_on_device = 0, shapes_on_device = 0, verts_on_device = 0;
    std::vector<double> pos3, rot4;  // packed staging copies of the padded real3 / quaternion arrays
};

}  // namespace chrono
