// libear-b200 C++14 facade, part 2: loudspeaker layouts and ADM metadata structs.
// ref include/ear/layout.hpp:15-96, bs2051.hpp:8-12, metadata.hpp:13-172 (same names/members).
#pragma once
#include <algorithm>
#include <memory>

#include "types.hpp"

namespace ear {

  class Channel {
   public:
    Channel() = default;
    Channel(const std::string &name, PolarPosition polarPosition,
            compat::optional<PolarPosition> polarPositionNominal = compat::none,
            compat::optional<std::pair<double, double>> azimuthRange = compat::none,
            compat::optional<std::pair<double, double>> elevationRange = compat::none, bool isLfe = false)
        : _name(name),
          _polarPosition(polarPosition),
          _polarPositionNominal(polarPositionNominal),
          _azimuthRange(azimuthRange),
          _elevationRange(elevationRange),
          _isLfe(isLfe) {}

    const std::string &name() const { return _name; }
    PolarPosition polarPosition() const { return _polarPosition; }
    PolarPosition polarPositionNominal() const {
      return _polarPositionNominal ? _polarPositionNominal.get() : _polarPosition;
    }
    std::pair<double, double> azimuthRange() const {
      return _azimuthRange ? _azimuthRange.get() : std::make_pair(_polarPosition.azimuth, _polarPosition.azimuth);
    }
    std::pair<double, double> elevationRange() const {
      return _elevationRange ? _elevationRange.get()
                             : std::make_pair(_polarPosition.elevation, _polarPosition.elevation);
    }
    bool isLfe() const { return _isLfe; }

    void name(const std::string &name) { _name = name; }
    void polarPosition(PolarPosition polarPosition) { _polarPosition = polarPosition; }
    void polarPositionNominal(const compat::optional<PolarPosition> &p) { _polarPositionNominal = p; }
    void azimuthRange(const compat::optional<std::pair<double, double>> &r) { _azimuthRange = r; }
    void elevationRange(const compat::optional<std::pair<double, double>> &r) { _elevationRange = r; }
    void isLfe(bool isLfe) { _isLfe = isLfe; }

   private:
    std::string _name;
    PolarPosition _polarPosition;
    compat::optional<PolarPosition> _polarPositionNominal;
    compat::optional<std::pair<double, double>> _azimuthRange;
    compat::optional<std::pair<double, double>> _elevationRange;
    bool _isLfe = false;
  };

  class Layout {
   public:
    Layout(std::string name = "", std::vector<Channel> channels = std::vector<Channel>(),
           compat::optional<Screen> screen = getDefaultScreen())
        : _name(std::move(name)), _channels(std::move(channels)), _screen(screen) {}

    std::string name() const { return _name; }
    std::vector<Channel> &channels() { return _channels; }
    const std::vector<Channel> &channels() const { return _channels; }
    compat::optional<Screen> screen() const { return _screen; }
    void name(std::string name) { _name = std::move(name); }
    void screen(compat::optional<Screen> screen) { _screen = screen; }

    Layout withoutLfe() const {
      Layout out(_name, {}, _screen);
      for (const Channel &c : _channels)
        if (!c.isLfe()) out._channels.push_back(c);
      return out;
    }
    std::vector<bool> isLfe() const {
      std::vector<bool> v;
      for (const Channel &c : _channels) v.push_back(c.isLfe());
      return v;
    }
    std::vector<std::string> channelNames() const {
      std::vector<std::string> v;
      for (const Channel &c : _channels) v.push_back(c.name());
      return v;
    }
    Channel channelWithName(const std::string &name) const {
      for (const Channel &c : _channels)
        if (c.name() == name) return c;
      throw invalid_argument("no channel named " + name);
    }
    compat::optional<int> indexForName(const std::string &name) const {
      for (size_t i = 0; i < _channels.size(); i++)
        if (_channels[i].name() == name) return (int)i;
      return compat::none;
    }
    std::vector<PolarPosition> positions() const {
      std::vector<PolarPosition> v;
      for (const Channel &c : _channels) v.push_back(c.polarPosition());
      return v;
    }
    std::vector<PolarPosition> nominalPositions() const {
      std::vector<PolarPosition> v;
      for (const Channel &c : _channels) v.push_back(c.polarPositionNominal());
      return v;
    }

   private:
    std::string _name;
    std::vector<Channel> _channels;
    compat::optional<Screen> _screen;
  };

  namespace b200 {
    struct LayoutDeleter {
      void operator()(earb200_layout *l) const { earb200_layout_destroy(l); }
    };
    using LayoutHandle = std::unique_ptr<earb200_layout, LayoutDeleter>;

    // C-ABI layout table -> ear::Layout
    inline Layout fromHandle(const std::string &name, const earb200_layout *h) {
      std::vector<Channel> channels;
      const int n = earb200_layout_num_channels(h);
      for (int i = 0; i < n; i++) {
        double pos[3], nom[3], az[2], el[2];
        check(earb200_layout_channel_position(h, i, pos, nom));
        check(earb200_layout_channel_ranges(h, i, az, el));
        channels.emplace_back(earb200_layout_channel_name(h, i), PolarPosition(pos[0], pos[1], pos[2]),
                              PolarPosition(nom[0], nom[1], nom[2]), std::make_pair(az[0], az[1]),
                              std::make_pair(el[0], el[1]), earb200_layout_channel_is_lfe(h, i) != 0);
      }
      return Layout(name, channels);
    }

    // ear::Layout -> C-ABI layout: the channels exactly as the Layout holds them.  A BS.2051
    // name selects the tabulated hull, any other name makes the panner triangulate the nominal
    // positions itself (ref src/common/point_source_panner.cpp:484-487); the panner range-checks
    // the screen loudspeakers as libear does (:560-578)
    inline LayoutHandle toHandle(const Layout &layout) {
      const std::vector<Channel> &ch = layout.channels();
      std::vector<const char *> names;
      std::vector<double> pos, nominal, az, el;
      std::vector<int32_t> lfe;
      for (const Channel &c : ch) {
        names.push_back(c.name().c_str());
        const PolarPosition p = c.polarPosition(), q = c.polarPositionNominal();
        pos.insert(pos.end(), {p.azimuth, p.elevation, p.distance});
        nominal.insert(nominal.end(), {q.azimuth, q.elevation, q.distance});
        az.insert(az.end(), {c.azimuthRange().first, c.azimuthRange().second});
        el.insert(el.end(), {c.elevationRange().first, c.elevationRange().second});
        lfe.push_back(c.isLfe() ? 1 : 0);
      }
      earb200_layout *h = nullptr;
      check(earb200_layout_create_custom(layout.name().c_str(), (int)ch.size(), names.data(), pos.data(),
                                         nominal.data(), az.data(), el.data(), lfe.data(), &h));
      return LayoutHandle(h);
    }
  }  // namespace b200

  // ref include/ear/bs2051.hpp:8-12, src/bs2051.cpp:13-22
  inline const std::vector<Layout> &loadLayouts() {
    static const std::vector<Layout> layouts = [] {
      std::vector<Layout> v;
      for (int i = 0; i < earb200_num_layouts(); i++) {
        const std::string name = earb200_layout_name_at(i);
        earb200_layout *h = nullptr;
        b200::check(earb200_layout_create(name.c_str(), &h));
        b200::LayoutHandle guard(h);
        v.push_back(b200::fromHandle(name, h));
      }
      return v;
    }();
    return layouts;
  }
  inline Layout getLayout(const std::string &name) {
    for (const Layout &l : loadLayouts())
      if (l.name() == name) return l;
    throw unknown_layout(name);
  }

  // ------------------------------------------------------------------- metadata
  // typeDefinition == DirectSpeakers (ref include/ear/metadata.hpp:13-72)
  struct ScreenEdgeLock {
    compat::optional<std::string> horizontal;
    compat::optional<std::string> vertical;
  };
  struct PolarSpeakerPosition {
    PolarSpeakerPosition(double az = 0.0, double el = 0.0, double dist = 1.0)
        : azimuth(az), elevation(el), distance(dist) {}
    double azimuth;
    compat::optional<double> azimuthMin, azimuthMax;
    double elevation;
    compat::optional<double> elevationMin, elevationMax;
    double distance;
    compat::optional<double> distanceMin, distanceMax;
    ScreenEdgeLock screenEdgeLock;
  };
  struct CartesianSpeakerPosition {
    CartesianSpeakerPosition(double X = 0.0, double Y = 1.0, double Z = 0.0) : X(X), Y(Y), Z(Z) {}
    double X;
    compat::optional<double> XMin, XMax;
    double Y;
    compat::optional<double> YMin, YMax;
    double Z;
    compat::optional<double> ZMin, ZMax;
    ScreenEdgeLock screenEdgeLock;
  };
  using SpeakerPosition = compat::variant2<PolarSpeakerPosition, CartesianSpeakerPosition>;
  struct ChannelFrequency {
    compat::optional<double> lowPass = compat::none;
    compat::optional<double> highPass = compat::none;
  };
  struct DirectSpeakersTypeMetadata {
    std::vector<std::string> speakerLabels = {};
    SpeakerPosition position = PolarSpeakerPosition();
    ChannelFrequency channelFrequency = {};
    compat::optional<std::string> audioPackFormatID = compat::none;
  };

  // typeDefinition == Objects (ref include/ear/metadata.hpp:76-145)
  struct ChannelLock {
    ChannelLock(bool flag = false, compat::optional<double> maxDistance = compat::none)
        : flag(flag), maxDistance(maxDistance) {}
    bool flag;
    compat::optional<double> maxDistance;
  };
  struct PolarObjectDivergence {
    PolarObjectDivergence(double divergence = 0.0, double azimuthRange = 45.0)
        : divergence(divergence), azimuthRange(azimuthRange) {}
    double divergence, azimuthRange;
  };
  struct CartesianObjectDivergence {
    CartesianObjectDivergence(double divergence = 0.0, double positionRange = 0.0)
        : divergence(divergence), positionRange(positionRange) {}
    double divergence, positionRange;
  };
  using ObjectDivergence = compat::variant2<PolarObjectDivergence, CartesianObjectDivergence>;
  struct PolarExclusionZone {
    float minAzimuth, maxAzimuth, minElevation, maxElevation, minDistance, maxDistance;
    std::string label;
  };
  struct CartesianExclusionZone {
    float minX, maxX, minY, maxY, minZ, maxZ;
    std::string label;
  };
  using ExclusionZone = compat::variant2<PolarExclusionZone, CartesianExclusionZone>;
  struct ZoneExclusion {
    std::vector<ExclusionZone> zones;
  };
  struct ObjectsTypeMetadata {
    Position position = {};
    double width = 0.0, height = 0.0, depth = 0.0;
    bool cartesian = false;
    double gain = 1.0, diffuse = 0.0;
    ChannelLock channelLock = {};
    ObjectDivergence objectDivergence = {};
    ZoneExclusion zoneExclusion = {};
    bool screenRef = false;
    Screen referenceScreen = getDefaultScreen();
  };

  // typeDefinition == HOA (ref include/ear/metadata.hpp:160-172)
  struct HOATypeMetadata {
    std::vector<int> orders;
    std::vector<int> degrees;
    std::string normalization = std::string("SN3D");
    double nfcRefDist = 0.0;
    bool screenRef = false;
    Screen referenceScreen = getDefaultScreen();
  };

}  // namespace ear
