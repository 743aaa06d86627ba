// libear-b200 C++14 facade, part 5: Cartesian <-> polar metadata conversion.
// ref include/ear/conversion.hpp:6-71 (same names and signatures), src/conversion.cpp.
// The single-object functions run the batched device pre-pass (earb200_conversion_*) with
// n = 1; the *Batch functions are the form a renderer should use for whole ADM blocks.
#pragma once
#include <memory>
#include <tuple>

#include "layout.hpp"
#include "types.hpp"

namespace ear {
  namespace conversion {
    /// extent parameters without an entire ObjectsTypeMetadata (ref conversion.hpp:8-12)
    struct ExtentParams {
      double width;
      double height;
      double depth;
    };

    namespace detail {
      struct Deleter {
        void operator()(earb200_conversion *c) const { earb200_conversion_destroy(c); }
      };
      // one converter per calling thread, on the thread's default device
      inline earb200_conversion *handle() {
        static thread_local std::unique_ptr<earb200_conversion, Deleter> h;
        static thread_local int device = -1;
        if (!h || device != b200::default_device()) {
          earb200_conversion *raw = nullptr;
          b200::check(earb200_conversion_create(b200::default_device(), &raw));
          h.reset(raw);
          device = b200::default_device();
        }
        return h.get();
      }
    }  // namespace detail

    // ---- batched forms (extension): pos / extent are [n][3] row-major, extent nullable
    inline void cartToPolarBatch(size_t n, const double *pos, const double *extent, double *pos_out,
                                 double *extent_out) {
      b200::check(earb200_conversion_cart_to_polar_host(detail::handle(), (int64_t)n, pos, extent, pos_out,
                                                        extent_out));
    }
    inline void polarToCartBatch(size_t n, const double *pos, const double *extent, double *pos_out,
                                 double *extent_out) {
      b200::check(earb200_conversion_polar_to_cart_host(detail::handle(), (int64_t)n, pos, extent, pos_out,
                                                        extent_out));
    }

    // ---- libear's functions
    inline PolarPosition pointCartToPolar(const CartesianPosition &pos) {
      const double in[3] = {pos.X, pos.Y, pos.Z};
      double out[3];
      cartToPolarBatch(1, in, nullptr, out, nullptr);
      return PolarPosition(out[0], out[1], out[2]);
    }
    inline CartesianPosition pointPolarToCart(const PolarPosition &pos) {
      const double in[3] = {pos.azimuth, pos.elevation, pos.distance};
      double out[3];
      polarToCartBatch(1, in, nullptr, out, nullptr);
      return CartesianPosition(out[0], out[1], out[2]);
    }
    inline std::pair<PolarPosition, ExtentParams> extentCartToPolar(const CartesianPosition &pos,
                                                                    const ExtentParams &extent) {
      const double in[3] = {pos.X, pos.Y, pos.Z}, ein[3] = {extent.width, extent.height, extent.depth};
      double out[3], eout[3];
      cartToPolarBatch(1, in, ein, out, eout);
      return {PolarPosition(out[0], out[1], out[2]), ExtentParams{eout[0], eout[1], eout[2]}};
    }
    inline std::pair<CartesianPosition, ExtentParams> extentPolarToCart(const PolarPosition &pos,
                                                                        const ExtentParams &extent) {
      const double in[3] = {pos.azimuth, pos.elevation, pos.distance},
                   ein[3] = {extent.width, extent.height, extent.depth};
      double out[3], eout[3];
      polarToCartBatch(1, in, ein, out, eout);
      return {CartesianPosition(out[0], out[1], out[2]), ExtentParams{eout[0], eout[1], eout[2]}};
    }

    /// in-place conversion of Objects metadata to polar; the type held by `position` decides,
    /// the cartesian flag is ignored and cleared (ref conversion.cpp:271-290)
    inline void toPolar(ObjectsTypeMetadata &otm) {
      otm.cartesian = otm.position.which() == 0;
      if (otm.cartesian) {
        PolarPosition polar_pos;
        ExtentParams polar_extent;
        std::tie(polar_pos, polar_extent) = extentCartToPolar(compat::get<CartesianPosition>(otm.position),
                                                              ExtentParams{otm.width, otm.height, otm.depth});
        otm.cartesian = false;
        otm.position = polar_pos;
        otm.width = polar_extent.width;
        otm.height = polar_extent.height;
        otm.depth = polar_extent.depth;
      }
    }
    /// in-place conversion of Objects metadata to Cartesian (ref conversion.cpp:292-311)
    inline void toCartesian(ObjectsTypeMetadata &otm) {
      otm.cartesian = otm.position.which() == 0;
      if (!otm.cartesian) {
        CartesianPosition cart_pos;
        ExtentParams cart_extent;
        std::tie(cart_pos, cart_extent) = extentPolarToCart(compat::get<PolarPosition>(otm.position),
                                                            ExtentParams{otm.width, otm.height, otm.depth});
        otm.cartesian = true;
        otm.position = cart_pos;
        otm.width = cart_extent.width;
        otm.height = cart_extent.height;
        otm.depth = cart_extent.depth;
      }
    }

    /// extension: toPolar over a whole block of objects in one device pass
    inline void toPolar(std::vector<ObjectsTypeMetadata> &block) {
      std::vector<size_t> idx;
      std::vector<double> pos, ext;
      for (size_t i = 0; i < block.size(); i++) {
        if (block[i].position.which() != 0) {
          block[i].cartesian = false;
          continue;
        }
        const CartesianPosition &c = compat::get<CartesianPosition>(block[i].position);
        idx.push_back(i);
        pos.insert(pos.end(), {c.X, c.Y, c.Z});
        ext.insert(ext.end(), {block[i].width, block[i].height, block[i].depth});
      }
      if (idx.empty()) return;
      std::vector<double> pos_out(pos.size()), ext_out(ext.size());
      cartToPolarBatch(idx.size(), pos.data(), ext.data(), pos_out.data(), ext_out.data());
      for (size_t k = 0; k < idx.size(); k++) {
        ObjectsTypeMetadata &otm = block[idx[k]];
        otm.cartesian = false;
        otm.position = PolarPosition(pos_out[3 * k], pos_out[3 * k + 1], pos_out[3 * k + 2]);
        otm.width = ext_out[3 * k];
        otm.height = ext_out[3 * k + 1];
        otm.depth = ext_out[3 * k + 2];
      }
    }
  }  // namespace conversion
}  // namespace ear
