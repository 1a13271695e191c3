/* stub: the mock Mitsuba of oracle/ref_mock.h stands in for <mitsuba/render/phase.h> (oracle/Makefile, _ref/refprobe_pm) */
