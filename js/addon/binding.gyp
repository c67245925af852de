{
  "targets": [{
    "target_name": "hlsmap_napi",
    "sources": ["hlsmap_napi.c"],
    "include_dirs": ["../../include"],
    "libraries": ["-L<(module_root_dir)/../../hlsjs-p2p-wrapper_b200/csrc", "-lhlsmap",
                  "-Wl,-rpath,<(module_root_dir)/../../hlsjs-p2p-wrapper_b200/csrc"]
  }]
}
