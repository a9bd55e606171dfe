// oracle/qt_shim_app/devbase_skel.h — TEST INFRASTRUCTURE ONLY.
// Class SKELETON of DeviceInterfaceBase (pebblelib/deviceinterfacebase.h:17-170): just the members and the
// method declarations that the normalizeIQ bodies (pebblelib/deviceinterfacebase.cpp:532-908) use.
// deviceinterfacebase.cpp as a whole needs QSettings/QThread/audio back ends and cannot be compiled here, so
// oracle/build_ref.sh step 5 cuts exactly those function bodies out of the reference file into a scratch
// translation unit that includes this header. No arithmetic is stated here.
#pragma once
#include "device_interfaces.h"

class DeviceInterfaceBase : public DeviceInterface {
public:
    DeviceInterfaceBase() : m_userIQGain(1.0), m_normalizeIQGain(1.0), m_iqOrder(IQO_IQ) {}
    ~DeviceInterfaceBase() {}
    bool initialize(CB_ProcessIQData, CB_ProcessBandscopeData, CB_ProcessAudioData, quint16) { return true; }
    bool command(StandardCommands, QVariant) { return true; }
    QVariant get(StandardKeys, QVariant = 0) { return QVariant(); }
    bool set(StandardKeys, QVariant, QVariant = 0) { return true; }

    // deviceinterfacebase.h:105-117
    void normalizeIQ(CPX *cpx, float I, float Q);
    void normalizeIQ(CPX *cpx, qint16 I, qint16 Q);
    void normalizeIQ(CPX *cpx, quint8 I, quint8 Q);
    void normalizeIQ(CPX *cpx, qint8 I, qint8 Q);
    void normalizeIQ(CPX *cpx, CPX iq);
    void normalizeIQ(CPX *_out, CPX8 *_in, quint32 _numSamples, bool _reverse = false);
    void normalizeIQ(CPX *_out, CPXU8 *_in, quint32 _numSamples, bool _reverse = false);
    void normalizeIQ(CPX *_out, CPX16 *_in, quint32 _numSamples, bool _reverse = false);
    void normalizeIQ(CPX *_out, CPX *_in, quint32 _numSamples, bool _reverse = false);
    void normalizeIQ(CPX *_out, CPXFLOAT *_in, quint32 _numSamples, bool _reverse = false);
    void normalizeIQ(CPX *_out, short *_inI, short *_inQ, quint32 _numSamples, bool _reverse = false);

    // deviceinterfacebase.h:73-76
    double m_userIQGain;
    double m_normalizeIQGain;
    IQOrder m_iqOrder;
};
