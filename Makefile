# Build of the hola service, its load generator, the C-ABI library and the C++ unit tests.
#
# The analogue of the reference's `mvn package` producing target/hola-swarm.jar
# (/root/reference/pom.xml:215-241): one self-contained executable, build/hola.
#
#   make            optimised build: build/hola build/hola-load build/hola-coldstart build/hola-null build/libhola.so build/hola_tests
#   make static     build/hola-static, fully static (for the FROM-scratch container image)
#   make asan       AddressSanitizer + UBSan variant under build/asan/
#   make tsan       ThreadSanitizer variant under build/tsan/
#   make test       run the C++ unit tests
#   make check      C++ tests + pytest + the black-box tests under ASan/UBSan and TSan
#   make cov        gcov-instrumented variant under build/cov/ (see tools/coverage.sh)
#   make clean

CXX      ?= g++
VARIANT  ?= release
BUILD    := build$(if $(filter release,$(VARIANT)),,/$(VARIANT))

CXXFLAGS_COMMON := -std=c++17 -Wall -Wextra -Wshadow -Wno-missing-field-initializers -Isrc -Iinclude -pthread -fPIC -MMD -MP \
                   -DHOLA_VERSION='"0.0.1-SNAPSHOT"'
CXXFLAGS_release := -O2 -g -DNDEBUG
CXXFLAGS_asan    := -O1 -g -fsanitize=address,undefined -fno-omit-frame-pointer
CXXFLAGS_tsan    := -O1 -g -fsanitize=thread -fno-omit-frame-pointer
CXXFLAGS_cov     := -O0 -g --coverage
LDFLAGS_release  :=
LDFLAGS_asan     := -fsanitize=address,undefined
LDFLAGS_tsan     := -fsanitize=thread
LDFLAGS_cov      := --coverage

CXXFLAGS := $(CXXFLAGS_COMMON) $(CXXFLAGS_$(VARIANT)) $(EXTRA_CXXFLAGS)
LDFLAGS  := -pthread $(LDFLAGS_$(VARIANT)) $(EXTRA_LDFLAGS)

CORE_SRC := $(wildcard src/core/*.cc)
NET_SRC  := $(wildcard src/net/*.cc)
APP_SRC  := $(filter-out src/app/main.cc,$(wildcard src/app/*.cc))
LIB_SRC  := $(CORE_SRC) $(NET_SRC) $(APP_SRC)
LIB_OBJ  := $(patsubst src/%.cc,$(BUILD)/obj/%.o,$(LIB_SRC))
TEST_SRC := $(wildcard tests/cpp/*.cc)
TEST_OBJ := $(patsubst tests/cpp/%.cc,$(BUILD)/obj/tests/%.o,$(TEST_SRC))

TARGETS := $(BUILD)/hola $(BUILD)/hola-load $(BUILD)/hola-coldstart $(BUILD)/hola-null $(BUILD)/libhola.so $(BUILD)/hola_tests

.PHONY: all static asan tsan cov test check clean
all: $(TARGETS)

$(BUILD)/obj/%.o: src/%.cc
	@mkdir -p $(dir $@)
	$(CXX) $(CXXFLAGS) -c $< -o $@

$(BUILD)/obj/tests/%.o: tests/cpp/%.cc
	@mkdir -p $(dir $@)
	$(CXX) $(CXXFLAGS) -Itests/cpp -c $< -o $@

$(BUILD)/libhola_core.a: $(LIB_OBJ)
	@rm -f $@
	ar rcs $@ $^

$(BUILD)/hola: $(BUILD)/obj/app/main.o $(BUILD)/libhola_core.a
	$(CXX) -o $@ $^ $(LDFLAGS)

$(BUILD)/hola-load: $(BUILD)/obj/tools/hola_load.o $(BUILD)/libhola_core.a
	$(CXX) -o $@ $^ $(LDFLAGS)

$(BUILD)/hola-coldstart: $(BUILD)/obj/tools/hola_coldstart.o $(BUILD)/libhola_core.a
	$(CXX) -o $@ $^ $(LDFLAGS)

$(BUILD)/hola-null: $(BUILD)/obj/tools/null_server.o $(BUILD)/libhola_core.a
	$(CXX) -o $@ $^ $(LDFLAGS)

$(BUILD)/libhola.so: $(BUILD)/obj/tools/capi.o $(BUILD)/libhola_core.a
	$(CXX) -shared -o $@ $^ $(LDFLAGS)

$(BUILD)/hola_tests: $(TEST_OBJ) $(BUILD)/libhola_core.a
	$(CXX) -o $@ $^ $(LDFLAGS)

static: $(BUILD)/hola-static
$(BUILD)/hola-static: $(BUILD)/obj/app/main.o $(BUILD)/libhola_core.a
	$(CXX) -static -o $@ $^ $(LDFLAGS)

# The sanitizer runtimes live next to the distribution compiler; a relocated toolchain named by
# $CXX may not find them, so these variants prefer /usr/bin/g++ when it exists.
SAN_CXX := $(firstword $(wildcard /usr/bin/g++) $(CXX))
asan:
	$(MAKE) VARIANT=asan CXX=$(SAN_CXX)
tsan:
	$(MAKE) VARIANT=tsan CXX=$(SAN_CXX)
# gcov instrumentation under build/cov/; see tools/coverage.sh
cov:
	$(MAKE) VARIANT=cov CXX=$(SAN_CXX)

test: $(BUILD)/hola_tests
	$(BUILD)/hola_tests

# Everything: C++ unit tests, the Python suite, then the black-box part again under both sanitizers.
BLACKBOX := tests/test_contract.py tests/test_server_core.py tests/test_chaining.py tests/test_tracing.py \
            tests/test_stream.py tests/test_dns.py tests/test_fuzz.py tests/test_cli.py
check: all
	$(BUILD)/hola_tests
	python -m pytest tests -q
	python -m pytest $(BLACKBOX) -q --io-backend epoll
	python -m pytest $(BLACKBOX) -q --variant asan
	python -m pytest $(BLACKBOX) -q --variant tsan

clean:
	rm -rf build

-include $(LIB_OBJ:.o=.d) $(TEST_OBJ:.o=.d) $(BUILD)/obj/app/main.d $(BUILD)/obj/tools/hola_load.d $(BUILD)/obj/tools/hola_coldstart.d $(BUILD)/obj/tools/capi.d $(BUILD)/obj/tools/null_server.d
