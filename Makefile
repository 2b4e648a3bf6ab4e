.POSIX:
.SUFFIXES:

# The reference's wrapper targets (Makefile:31-59 there): `make standalone`, `make python`, `make clean` drive CMake;
# `make inplace` is the in-tree build the tests, bench.py and __graft_entry__.build() use (lib/ under the package).
CMAKE      = cmake
BUILD_DIR  = build
BUILD_TYPE = Release
BUILD_STANDALONE_DIR = $(BUILD_DIR)/standalone
BUILD_PYTHON_DIR     = $(BUILD_DIR)/python
PACKAGE    = chaos-ray-tracing-course-2025_b200

.PHONY: standalone python inplace clean

standalone:
	mkdir -p $(BUILD_STANDALONE_DIR)
	$(CMAKE) -S . -B $(BUILD_STANDALONE_DIR) -DCMAKE_BUILD_TYPE=$(BUILD_TYPE) -DBUILD_STANDALONE=ON -DBUILD_PYTHON=OFF
	$(CMAKE) --build $(BUILD_STANDALONE_DIR) -j
	@echo "> Run with ./$(BUILD_STANDALONE_DIR)/crt_renderer <scene_file> <output_file>"

python:
	mkdir -p $(BUILD_PYTHON_DIR)
	$(CMAKE) -S . -B $(BUILD_PYTHON_DIR) -DCMAKE_BUILD_TYPE=$(BUILD_TYPE) -DBUILD_STANDALONE=OFF -DBUILD_PYTHON=ON
	$(CMAKE) --build $(BUILD_PYTHON_DIR) --target _crt -j
	@echo "> You can import the _crt module from ./$(BUILD_PYTHON_DIR)"

inplace:
	$(MAKE) -C $(PACKAGE)

clean:
	rm -rf $(BUILD_DIR)
	$(MAKE) -C $(PACKAGE) clean
