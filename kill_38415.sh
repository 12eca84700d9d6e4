ssh test@127.0.0.1 "kill -9 2248"
ssh test@127.0.0.1 "kill -9 2249"
ssh test@127.0.0.1 "kill -9 2250"
ssh test@127.0.0.1 "kill -9 2251"
kill -9 2109
