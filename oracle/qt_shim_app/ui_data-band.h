// oracle/qt_shim_app — TEST INFRASTRUCTURE ONLY. Stand-in for the uic-generated header demod.h:13 includes.
#pragma once
class QWidget;
namespace Ui { class dataBand { public: void setupUi(QWidget *) {} }; }
