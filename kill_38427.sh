ssh test@127.0.0.1 "kill -9 2134"
ssh test@127.0.0.1 "kill -9 2135"
kill -9 2109
