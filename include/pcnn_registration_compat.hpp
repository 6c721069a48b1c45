// pcnn_registration_compat.hpp -- the ICP half of the reference's `Registration` class on top of
// the C ABI (include/pcnn.h), for callers written against
//   Homework9/hw9/include/registration.hpp:126-137  setICPparams(normal_bins, sampled_size,
//                                                   max_corres_dist, max_iter, loss_epsilon)
//   Homework9/hw9/src/registration.cpp:710-717      ICPpoint2plane(init_R, init_t, src, tar,
//                                                   normals_src, normals_tar, R, t)
//   Homework9/hw9/src/registration.cpp:862-869      ICPpoint2point(... same argument list ...)
//   Homework9/hw9/src/registration.cpp:165-178      transformCloudInplace(cloud, R, t)
//
// The reference's types come from Eigen and PCL, neither of which this repository may assume.
// The class is therefore a template over them; a reference build instantiates it with its own:
//
//   #include "pcnn_registration_compat.hpp"
//   using Registration = pcnn_compat::RegistrationT<Eigen::Matrix3f, Eigen::Vector3f,
//                                                   PointCloud, NormalCloud>;
//
// What the type parameters must offer (Eigen and PCL do):
//   Mat3   float& operator()(int r, int c)          Vec3   float& operator()(int i)
//   Cloud  size(), operator[](i) with .x .y .z      Normals size(), operator[](i) with
//                                                          .normal_x .normal_y .normal_z
//
// Behaviour kept as the reference has it:
//  - the initial pose applied to the source cloud is the OUTPUT pair R, t, not init_R / init_t
//    (registration.cpp:719-722, :871-875; compute() passes the same values for both, :1141-1151);
//  - max_corres_dist gates the SQUARED 1-NN distance (:783, :936);
//  - the loop ends early once |last_loss - loss| < loss_epsilon has held more than 15 times;
//  - R, t return T_total's rotation and translation (:856-857, :1008-1009).
// What differs: the normal-space sampling in front of the loop (:728-729, :880-881, PCL) is not
// done here -- the clouds handed in are the clouds the loop runs on, as with the reference's own
// bypass at :731-732 / :883-884; normal_bins and sampled_size are stored and ignored.  Errors
// throw std::runtime_error carrying pcnn_last_error().  There is no CPU fallback.
#ifndef PCNN_REGISTRATION_COMPAT_HPP
#define PCNN_REGISTRATION_COMPAT_HPP

#include <cstddef>
#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>

#include "pcnn.h"

namespace pcnn_compat {

// transformCloudInplace (registration.cpp:165-178): p <- R p + t, float32, in place.
template <class Cloud, class Mat3, class Vec3>
void transformCloudInplace(Cloud& cloud, const Mat3& R, const Vec3& t) {
    for (std::size_t i = 0; i < cloud.size(); i++) {
        const float x = cloud[i].x, y = cloud[i].y, z = cloud[i].z;
        cloud[i].x = ((R(0, 0) * x + R(0, 1) * y) + R(0, 2) * z) + t(0);
        cloud[i].y = ((R(1, 0) * x + R(1, 1) * y) + R(1, 2) * z) + t(1);
        cloud[i].z = ((R(2, 0) * x + R(2, 1) * y) + R(2, 2) * z) + t(2);
    }
}

template <class Mat3, class Vec3, class Cloud, class Normals>
class RegistrationT {
public:
    RegistrationT() = default;

    // registration.hpp:126-137
    void setICPparams(const int normal_bins, const std::size_t sampled_size, const float max_corres_dist,
                      const std::size_t max_iter, const float loss_epsilon) {
        m_ICP_normal_bins = normal_bins;
        m_ICP_sampled_size = sampled_size;
        m_ICP_max_corres_dist = max_corres_dist;
        m_ICP_max_iter = max_iter;
        m_ICP_loss_epsilon = loss_epsilon;
    }
    void setDevice(int device) { m_device = device; }

    // registration.cpp:862-1011
    void ICPpoint2point(const Mat3& init_R, const Vec3& init_t, const Cloud& input_cloud_src,
                        const Cloud& input_cloud_tar, const Normals& input_normals_src,
                        const Normals& input_normals_tar, Mat3& R, Vec3& t) {
        (void)init_R; (void)init_t; (void)input_normals_src;
        run(0, input_cloud_src, input_cloud_tar, input_normals_tar, R, t);
    }
    // registration.cpp:710-859
    void ICPpoint2plane(const Mat3& init_R, const Vec3& init_t, const Cloud& input_cloud_src,
                        const Cloud& input_cloud_tar, const Normals& input_normals_src,
                        const Normals& input_normals_tar, Mat3& R, Vec3& t) {
        (void)init_R; (void)init_t; (void)input_normals_src;
        run(1, input_cloud_src, input_cloud_tar, input_normals_tar, R, t);
    }

    // what the last call did (the reference only prints these, :822-826, :943-946)
    const pcnn_icp_result& lastResult() const { return m_last; }

private:
    void run(int mode, const Cloud& src, const Cloud& tar, const Normals& tar_normals, Mat3& R, Vec3& t) {
        std::vector<float> s(3 * src.size()), d(3 * tar.size()), n;
        for (std::size_t i = 0; i < src.size(); i++) { s[3 * i] = src[i].x; s[3 * i + 1] = src[i].y; s[3 * i + 2] = src[i].z; }
        for (std::size_t i = 0; i < tar.size(); i++) { d[3 * i] = tar[i].x; d[3 * i + 1] = tar[i].y; d[3 * i + 2] = tar[i].z; }
        if (mode == 1) {
            if (tar_normals.size() != tar.size()) throw std::runtime_error("ICPpoint2plane: one normal per target point is required");
            n.resize(3 * tar.size());
            for (std::size_t i = 0; i < tar.size(); i++) {
                n[3 * i] = tar_normals[i].normal_x; n[3 * i + 1] = tar_normals[i].normal_y; n[3 * i + 2] = tar_normals[i].normal_z;
            }
        }
        pcnn_index* index = nullptr;
        if (pcnn_index_build_host(&index, PCNN_KDTREE, d.data(), (int64_t)tar.size(), 3, 2 /* nanoflann leaf, :753 */, 0.0,
                                  m_device) != PCNN_OK)
            throw std::runtime_error(std::string("Registration: ") + pcnn_last_error());
        float T0[16] = {R(0, 0), R(0, 1), R(0, 2), t(0), R(1, 0), R(1, 1), R(1, 2), t(1),
                        R(2, 0), R(2, 1), R(2, 2), t(2), 0.f, 0.f, 0.f, 1.f};
        pcnn_icp_params prm;
        prm.max_iter = (int32_t)m_ICP_max_iter;
        prm.max_corres_dist = m_ICP_max_corres_dist;
        prm.loss_epsilon = m_ICP_loss_epsilon;
        prm.mode = mode;
        const int rc = pcnn_icp_host(index, mode == 1 ? n.data() : nullptr, s.data(), (int64_t)src.size(), 3, T0, &prm, &m_last);
        pcnn_index_destroy(index);
        if (rc != PCNN_OK) throw std::runtime_error(std::string("Registration: ") + pcnn_last_error());
        for (int r = 0; r < 3; r++) {
            for (int c = 0; c < 3; c++) R(r, c) = m_last.T[4 * r + c];
            t(r) = m_last.T[4 * r + 3];
        }
    }

    int m_ICP_normal_bins = 10;              // main.cpp:88-95
    std::size_t m_ICP_sampled_size = 4000;
    float m_ICP_max_corres_dist = 1.f;
    std::size_t m_ICP_max_iter = 800;
    float m_ICP_loss_epsilon = 1e-8f;
    int m_device = 0;
    pcnn_icp_result m_last{};
};

}  // namespace pcnn_compat

#endif  // PCNN_REGISTRATION_COMPAT_HPP
