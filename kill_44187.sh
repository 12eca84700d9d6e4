ssh test@127.0.0.1 "kill -9 1894"
ssh test@127.0.0.1 "kill -9 1895"
kill -9 1591
