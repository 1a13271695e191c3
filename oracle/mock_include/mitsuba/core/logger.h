/* stub: the mock Mitsuba of oracle/ref_mock.h stands in for <mitsuba/core/logger.h> (oracle/Makefile, _ref/refprobe_lc) */
