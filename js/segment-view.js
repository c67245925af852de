// Same module path and default export as lib/integration/mapping/segment-view.js
module.exports = require('./mapping').SegmentView;
module.exports.default = module.exports;
